"""-m gpu: the data-parallel optimiser over peer memory (csrc/peer.cu, optim.PeerAdam).

world 1 (any box): PeerAdam degenerates to ScaledAdam -- the pull is the identity, the barrier meets itself -- and must
produce the same bits.  world 2 (needs two GPUs; skipped otherwise): two processes, one per GPU, peer arenas opened over
CUDA IPC; every rank must end with the fp16 table and the replicated tensors that ONE ScaledAdam fed the fp32 mean of
the two gradient sets produces, an overflow on one rank must skip the step on both, and sync_master() must complete the
sharded fp32 tensors."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

TABLE, TAIL = (50000, 2), [3072, 7168]          # table rows x level_dim; the two MLP weight tensors of the field


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _init(seed, dev):
    g = torch.Generator(device=dev).manual_seed(seed)
    return [torch.randn(TABLE, device=dev, generator=g) * 0.1] + [torch.randn(n, device=dev, generator=g) * 0.1 for n in TAIL]


def _grads(it, rank, dev, world):
    g = torch.Generator(device=dev).manual_seed(1000 * it + rank)
    out = [(torch.randn(TABLE, device=dev, generator=g) * 1e-3)] + [torch.randn(n, device=dev, generator=g) * 1e-3 for n in TAIL]
    if it == 4 and rank == 0:
        out[0][17, 1] = float("inf")            # one rank overflows: everybody skips
    if it == 6 and rank == world - 1:
        out[2][5] = float("nan")
    return out


def _expected(world, steps, dev):
    """One ScaledAdam stepping on the fp32 mean over ranks of the fp16-rounded scaled table gradients / fp32 tail gradients."""
    from focnerf_b200.optim import ScaledAdam
    ps = [torch.nn.Parameter(t.clone()) for t in _init(0, dev)]
    opt = ScaledAdam(ps, lr=1e-2, betas=(0.9, 0.99), eps=1e-15, growth_interval=3)
    for it in range(steps):
        scale = float(opt.scale.item())
        per_rank = [_grads(it, r, dev, world) for r in range(world)]
        tab = sum((g[0] * scale).half().float() for g in per_rank) * (1.0 / world)      # rank order, fp32 accumulation
        ps[0].grad = tab
        for k in (1, 2):
            acc = torch.zeros_like(ps[k])
            for g in per_rank:
                acc = acc + g[k] * scale
            ps[k].grad = acc * (1.0 / world)
        opt.step()
        opt.zero_grad()
    return ps, opt


def _run_rank(rank, world, port, fused, result):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), FOC_DP_FUSED=str(fused))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from focnerf_b200.optim import PeerAdam
    steps = 9
    ps = [torch.nn.Parameter(t.clone()) for t in _init(0, dev)]
    opt = PeerAdam(ps, table=ps[0], lr=1e-2, betas=(0.9, 0.99), eps=1e-15, growth_interval=3)
    assert opt.fused_exchange == bool(fused)
    for it in range(steps):
        scale = float(opt.scale.item())
        g = _grads(it, rank, dev, world)
        opt.grad16.copy_((g[0] * scale).half())
        ps[1].grad, ps[2].grad = g[1] * scale, g[2] * scale
        opt.step()
        opt.zero_grad()
    torch.cuda.synchronize()
    exp, eopt = _expected(world, steps, dev)
    ok = {}
    ok["no_timeout"] = not opt.arena.timed_out()
    ok["steps"] = int(opt.step_count.item()) == int(eopt.step_count.item()) == steps - 2
    ok["scale"] = float(opt.scale.item()) == float(eopt.scale.item())
    ok["shadow"] = torch.equal(opt.shadow16, exp[0].detach().half())
    ok["tail"] = all(torch.equal(ps[k].detach(), exp[k].detach()) for k in (1, 2))
    ok["grad_zeroed"] = float(opt.grad16.float().abs().max()) == 0.0
    opt.sync_master()
    ok["master"] = torch.equal(ps[0].detach(), exp[0].detach())
    ti = 0
    ok["moments"] = torch.equal(opt.exp_avg[ti], eopt.exp_avg[ti]) and torch.equal(opt.exp_avg_sq[ti], eopt.exp_avg_sq[ti])
    result[rank] = ok
    dist.barrier()
    opt.arena.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("fused", [1, 0])       # 1: exchange inside the sweep (foc_adam_step_dp, two barriers); 0: foc_dp_reduce + foc_adam_step_peers
@pytest.mark.parametrize("world", [1, 2])
def test_peer_adam_matches_scaled_adam_on_mean_gradient(world, fused):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    mgr = mp.Manager()
    result = mgr.dict()
    mp.spawn(_run_rank, args=(world, _free_port(), fused, result), nprocs=world, join=True)
    for r in range(world):
        assert all(result[r].values()), (r, dict(result[r]))


def _train_rank(rank, world, port, mode, result):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl" if mode == "nccl" else "gloo", rank=rank, world_size=world)
    import cases
    from focnerf_b200.trainer import Trainer
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid), device=dev, seed=0, use_graph=True, data_parallel=mode, lr=1e-2)
    losses = []
    for it in range(40):
        o, d = cases.camera_rays(1024, seed=100 + 2 * it + rank)          # every rank its own rays
        o, d = torch.from_numpy(o).to(dev), torch.from_numpy(d).to(dev)
        gt = scene.render_gt(o, d)
        losses.append(tr.train_step(o, d, gt.to(dev)).detach().clone())
        if tr.local_step % 16 == 0:
            tr.update_mean_count()
    torch.cuda.synchronize()
    loss = torch.stack([l.detach().float().reshape(()) for l in losses]).cpu()
    sh = tr.model.encoder._shadow16
    digest = torch.stack([sh.float().sum(), sh.float().abs().sum(), tr.model.sigma_net.weights.detach().sum(), tr.model.color_net.weights.detach().sum()]).cpu()
    result[rank] = (loss, digest, (not tr.optimizer.arena.timed_out()) if mode == "peer" else True)
    dist.barrier()
    tr._graph = None
    torch.cuda.synchronize()
    dist.destroy_process_group()


def test_data_parallel_trainer_peer_vs_nccl():
    """Two ranks, graph-captured steps: the replicas stay bit-identical under the peer exchange, and the loss curve matches
    the NCCL all-reduce + replicated sweep it replaces (fp16 ring sum there, fp32 rank-order sum here: close, not equal)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    mgr = mp.Manager()
    out = {}
    for mode in ("peer", "nccl"):
        result = mgr.dict()
        mp.spawn(_train_rank, args=(2, _free_port(), mode, result), nprocs=2, join=True)
        out[mode] = {r: result[r] for r in range(2)}
        assert out[mode][0][2] and out[mode][1][2]
        assert torch.equal(out[mode][0][1], out[mode][1][1]), (mode, out[mode][0][1], out[mode][1][1])     # replicas in sync
    lp, ln = out["peer"][0][0], out["nccl"][0][0]
    assert float(lp[-8:].mean()) < 0.6 * float(lp[:4].mean())                                              # it trains
    assert abs(float(lp[-8:].mean()) - float(ln[-8:].mean())) < 0.15 * float(ln[-8:].mean()), (lp[-8:], ln[-8:])


def test_render_finish_matches_torch_ops():
    """The fused render epilogue = the reference's separate elementwise ops (renderer.py:369-372), bit for bit."""
    from focnerf_b200 import peer
    g = torch.Generator(device="cuda").manual_seed(5)
    n = 1000                                                          # not a multiple of the 128-ray block
    image, depth, ws = torch.rand(n, 3, device="cuda", generator=g), torch.rand(n, device="cuda", generator=g) * 3, torch.rand(n, device="cuda", generator=g)
    ws[::7] = 0.0
    nears = torch.rand(n, device="cuda", generator=g) + 0.2
    fars = nears + torch.rand(n, device="cuda", generator=g) * 3 + 0.1
    for from_first in (False, True):
        buf, out = peer.local_frame(n, image.device)
        buf.fill_(-7.0)
        peer.render_finish(image, depth, ws, nears, fars, 0.75, from_first, out)
        d = depth + torch.where(ws > 0, nears, torch.zeros_like(nears)) * ws if from_first else depth
        want = torch.cat([image + (1 - ws).unsqueeze(-1) * 0.75, (torch.clamp(d - nears, min=0) / (fars - nears)).unsqueeze(-1), ws.unsqueeze(-1)], -1)
        assert torch.equal(buf, want)


def _render_rank(rank, world, port, result):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import cases
    from focnerf_b200.peer import FrameExchange
    from focnerf_b200.trainer import Trainer
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid), device=dev, seed=0)      # same seed: identical replicas
    n = 5000                                                                             # 40 blocks of 128 rays, the last one short
    o, d = cases.camera_rays(n, seed=9)
    o, d = torch.from_numpy(o).to(dev), torch.from_numpy(d).to(dev)
    ex = FrameExchange(n, device=dev)
    idx = ex.my_rays(device=dev)
    ok = {}
    for k in range(3):                                                                   # both frame buffers, and the first one again
        img, dep, ws = tr.render(o[idx], d[idx], exchange=ex)
        full = tr.render(o, d)
        torch.cuda.synchronize()
        ok[f"shape{k}"] = img.shape[0] == n
        for name, a, b in zip(("image", "depth", "ws"), (img, dep, ws), full):
            ok[f"{name}{k}"] = torch.equal(a.nan_to_num(nan=-1.0), b.nan_to_num(nan=-1.0)) or float((a - b).abs().max())   # 0/0 depth on rays that miss the box
    ok["no_timeout"] = not ex.arena.timed_out()
    ok["covered"] = bool((ws > 0).any())
    result[rank] = ok
    dist.barrier()
    ex.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2])
def test_frame_exchange_equals_single_gpu_render(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    mgr = mp.Manager()
    result = mgr.dict()
    mp.spawn(_render_rank, args=(world, _free_port(), result), nprocs=world, join=True)
    for r in range(world):
        assert all(v is True for v in result[r].values()), (r, dict(result[r]))
