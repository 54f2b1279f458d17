"""Host-side multi-GPU logic of the path: rays shard, parameters replicate.

* rendering: a static partition of the H*W rays into contiguous, 128-aligned tiles, one per rank, no collective on the
  data path; the finished tiles are gathered once;
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


def gather_tiles(tile, n_rays, group=None, align=128):
    """All ranks contribute their [tile_len, C] slice; returns the assembled [n_rays, C] tensor on every rank."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    per = ray_tile(n_rays, 0, world, align)[1]
    padded = tile.new_zeros((per,) + tuple(tile.shape[1:]))
    padded[:tile.shape[0]] = tile
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded, group=group)
    out = torch.cat(parts, 0)[:n_rays]
    s, e = ray_tile(n_rays, rank, world, align)
    assert tile.shape[0] == e - s
    return out
