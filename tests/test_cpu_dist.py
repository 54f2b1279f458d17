"""CPU, world_size 2 over gloo: the host-side N>1 logic (ray tiles, gradient averaging, tile gather, replica broadcast)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from focnerf_b200 import dist as fd
    n = 1000
    s, e = fd.ray_tile(n, rank, world)
    img = torch.arange(n * 3, dtype=torch.float32).view(n, 3)
    full = fd.gather_tiles(img[s:e].clone(), n)
    ok_gather = torch.equal(full, img)
    idx = fd.cyclic_tile(n, rank, world)
    frame, _ = fd.gather_cyclic(img[idx].clone(), n)
    ok_gather = ok_gather and torch.equal(frame, img) and idx.shape[0] == 512
    g32 = torch.full((5,), float(rank + 1))
    g16 = torch.full((7,), float(rank + 1), dtype=torch.float16)
    fd.all_reduce_mean([g32, g16])
    ok_mean = torch.allclose(g32, torch.full((5,), 1.5)) and torch.allclose(g16.float(), torch.full((7,), 1.5))
    p = torch.nn.Parameter(torch.full((4,), float(rank)))
    fd.broadcast_parameters([p])
    ok_bcast = bool((p.data == 0).all())
    out[rank] = (ok_gather, ok_mean, ok_bcast, (s, e))
    dist.destroy_process_group()


def test_world2_gloo():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert out[0][:3] == (True, True, True) and out[1][:3] == (True, True, True)
    assert out[0][3] == (0, 512) and out[1][3] == (512, 1000)


def test_ray_tiles_cover():
    from focnerf_b200.dist import ray_tile
    for n in (0, 1, 127, 128, 129, 640000, 307200):
        for world in (1, 2, 4, 8):
            pos = 0
            for r in range(world):
                s, e = ray_tile(n, r, world)
                assert s == pos and e >= s and (s % 128 == 0 or s == n)
                pos = e
            assert pos == n


def test_cyclic_tiles_cover():
    from focnerf_b200.dist import cyclic_tile
    for n in (1, 127, 128, 129, 1000, 640000):
        for world in (1, 2, 4, 8):
            seen = torch.zeros(n, dtype=torch.int32)
            lens = set()
            for r in range(world):
                idx = cyclic_tile(n, r, world)
                lens.add(idx.shape[0])
                seen[idx] += 1
                blk = idx[::128] // 128
                assert bool(((blk % world == r) | (idx[::128] == n - 1)).all())
            assert len(lens) == 1 and int(seen.min()) >= 1
            assert int((seen[:-1] > 1).sum()) == 0          # only the last ray may repeat (padding)
