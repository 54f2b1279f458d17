"""-m gpu, collected LAST on purpose (pytest -x: a failure here must not hide other tests).

foc_density_grid_update_noalloc (include/foc_b200.h): the density-grid update with its scratch in a static device array inside the
library -- for callers that cannot cudaMalloc (graph capture, serving loops).  Called through the C ABI (ctypes, raw pointers) and
compared with the existing foc_density_grid_update on the same inputs, on two calls in a row (the second smaller, so stale contents
of the shared array from the first call lie behind what it uses).  Same kernels and the same fixed summation order: grid and
bitfield must be bit-identical, the mean within 1e-6 relative (the tolerance class of the other float reductions in this suite)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

FOC_ERR_TOO_LARGE = -2
MAX_CELLS = 8 * 128 ** 3


def _inputs(n, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    grid = torch.rand(n, generator=g) * 30.0
    grid[torch.rand(n, generator=g) < 0.1] = -1.0            # untrained cells (mark_untrained_grid) stay out of the EMA
    tmp = torch.rand(n, generator=g) * 40.0
    tmp[torch.rand(n, generator=g) < 0.5] = -1.0             # cells not re-sampled by a partial update
    return grid.cuda(), tmp.cuda()


def _run_pair(lib, n, seed, decay=0.95, thresh=10.0):
    grid, tmp = _inputs(n, seed)
    ga, gb = grid.clone(), grid.clone()
    mean_a, mean_b = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
    bits_a = torch.zeros(n // 8, dtype=torch.uint8, device="cuda")
    bits_b = torch.full((n // 8,), 0xAA, dtype=torch.uint8, device="cuda")
    scratch = torch.empty(max(4, lib.foc_density_grid_update_scratch_bytes(n)), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert lib.foc_density_grid_update(ga.data_ptr(), tmp.data_ptr(), n, decay, thresh, mean_a.data_ptr(), bits_a.data_ptr(),
                                       scratch.data_ptr(), st) == 0
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.foc_density_grid_update_noalloc(gb.data_ptr(), tmp.data_ptr(), n, decay, thresh, mean_b.data_ptr(), bits_b.data_ptr(), st)
    assert rc == 0, rc
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before           # nothing came from torch's allocator either
    assert torch.equal(ga, gb), "EMA grid differs from foc_density_grid_update"
    assert torch.equal(bits_a, bits_b), "bitfield differs from foc_density_grid_update"
    a, b = float(mean_a), float(mean_b)
    assert a > 0 and abs(a - b) <= 1e-6 * abs(a), (a, b)
    # and against a plain restatement (renderer.py:522-530), so that both being wrong together would show
    valid = (grid >= 0) & (tmp >= 0)
    exp = torch.where(valid, torch.maximum(grid * decay, tmp), grid)
    assert torch.equal(gb, exp)
    exp_mean = float(exp.clamp(min=0).double().mean())
    assert abs(b - exp_mean) <= 1e-4 * exp_mean
    return b


def test_noalloc_density_grid_update_matches_existing_entry_point_twice():
    from focnerf_b200 import _lib
    lib = _lib.lib()
    m1 = _run_pair(lib, 2 * 128 ** 3, seed=1)                # two cascades: 2048 partial sums in the static array
    m2 = _run_pair(lib, 128 ** 3 // 4, seed=2)               # right after: 256 partial sums, the rest of the array is stale
    m3 = _run_pair(lib, 2 * 128 ** 3, seed=1)                # and the first call again gives the first answer
    assert m1 == m3 and m1 != m2


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    from focnerf_b200 import _lib
    lib = _lib.lib()
    n = MAX_CELLS + 8
    one = torch.zeros(8, device="cuda")                      # never touched: the call must return before launching anything
    rc = lib.foc_density_grid_update_noalloc(one.data_ptr(), one.data_ptr(), n, 0.95, 10.0, one.data_ptr(), one.data_ptr(),
                                             torch.cuda.current_stream().cuda_stream)
    assert rc == FOC_ERR_TOO_LARGE
    torch.cuda.synchronize()
    assert float(one.abs().sum()) == 0.0


def test_noalloc_accepts_exactly_the_maximum():
    from focnerf_b200 import _lib
    _run_pair(_lib.lib(), MAX_CELLS, seed=3)                 # 8 cascades of 128^3: all 8192 slots of the static array
