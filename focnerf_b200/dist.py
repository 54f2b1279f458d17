"""Host-side multi-GPU logic of the path: rays shard, parameters replicate.

* rendering: a static partition of the H*W rays into 128-ray blocks dealt round-robin to the ranks (`cyclic_tile`: a
  contiguous split leaves the ranks holding empty background strips idle -- measured 78 % of the samples of the bench
  frame in its top half), no collective on the data path; the finished tiles are gathered and un-dealt once per frame
  (`gather_cyclic`).  `ray_tile` / `gather_tiles` are the contiguous variant;
* training: data parallel -- every rank marches its own rays, the gradient tensors are averaged between backward and
  the optimiser sweep.
One process per GPU, `torch.distributed` (NCCL on GPUs; the same code runs on gloo/CPU tensors for the tests)."""
import torch
import torch.distributed as dist


def ray_tile(n_rays, rank, world, align=128):
    """[start, end) of `rank`'s contiguous ray tile: tiles are equal, rounded up to `align` rays, the last one short."""
    per = -(-n_rays // world)
    per = -(-per // align) * align
    start = min(rank * per, n_rays)
    return start, min(start + per, n_rays)


def all_reduce_mean(tensors, group=None):
    """In-place average of every tensor over the group (one collective per tensor: the table gradient is the only large
    one).  NCCL averages natively; gloo sums and the division is done here."""
    world = dist.get_world_size(group)
    if world == 1:
        return
    native_avg = dist.get_backend(group) == "nccl"
    for t in tensors:
        if native_avg:
            dist.all_reduce(t, op=dist.ReduceOp.AVG, group=group)
        else:
            acc = t.float() if t.dtype == torch.float16 else t          # gloo has no fp16 reduction
            dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
            acc /= world
            if acc is not t:
                t.copy_(acc)


def broadcast_parameters(params, src=0, group=None):
    for p in params:
        dist.broadcast(p.data, src, group=group)


def gather_tiles(tile, n_rays, group=None, align=128, out=None):
    """All ranks contribute their [tile_len, C] slice; returns the assembled [n_rays, C] tensor on every rank.
    One all_gather_into_tensor over equal (padded) tiles, straight into `out` (a [world * per, C] buffer the caller may
    keep across frames); the result is a view of it."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    per = ray_tile(n_rays, 0, world, align)[1]
    s, e = ray_tile(n_rays, rank, world, align)
    assert tile.shape[0] == e - s
    shape = (world * per,) + tuple(tile.shape[1:])
    if out is None or tuple(out.shape) != shape or out.dtype != tile.dtype or out.device != tile.device:
        out = tile.new_empty(shape)
    if tile.shape[0] == per and tile.is_contiguous():
        padded = tile
    else:
        padded = tile.new_zeros((per,) + tuple(tile.shape[1:]))
        padded[:tile.shape[0]] = tile
    if dist.get_backend(group) == "nccl":
        dist.all_gather_into_tensor(out, padded, group=group)
    else:                                                            # gloo: list form
        parts = list(out.view((world, per) + tuple(tile.shape[1:])).unbind(0))
        dist.all_gather(parts, padded, group=group)
    return out[:n_rays]


def cyclic_tile(n_rays, rank, world, block=128, device=None):
    """Ray indices of `rank` under a block-cyclic deal: block b (rays [b*block, (b+1)*block)) belongs to rank b % world.
    Every rank gets the same number of blocks; indices past the end repeat the last ray (rendered twice, dropped by
    gather_cyclic).  -> int64 [blocks_per_rank * block]."""
    nblk = -(-n_rays // block)
    per = -(-nblk // world)
    b = torch.arange(per, device=device) * world + rank
    idx = (b[:, None] * block + torch.arange(block, device=device)[None, :]).reshape(-1)
    return idx.clamp_(max=n_rays - 1)


def gather_cyclic(tile, n_rays, group=None, block=128, out=None):
    """Inverse of cyclic_tile: every rank contributes its [blocks_per_rank * block, C] tile, every rank gets the
    [n_rays, C] frame back in ray order.  One all_gather_into_tensor ([world, per, block, C]) + one transposing copy."""
    world = dist.get_world_size(group)
    nblk = -(-n_rays // block)
    per = -(-nblk // world)
    C = tuple(tile.shape[1:])
    assert tile.shape[0] == per * block
    shape = (world, per, block) + C
    if out is None or tuple(out.shape) != shape or out.dtype != tile.dtype or out.device != tile.device:
        out = tile.new_empty(shape)
    tile = tile.contiguous()
    if dist.get_backend(group) == "nccl":
        dist.all_gather_into_tensor(out, tile.view((1,) + shape[1:]), group=group)
    else:
        dist.all_gather(list(out.unbind(0)), tile.view(shape[1:]), group=group)
    frame = out.transpose(0, 1).reshape((per * world * block,) + C)
    return frame[:n_rays], out
