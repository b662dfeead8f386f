"""kst_objective_batch_noalloc: the entry point for callers that must not allocate device memory while
they run (their memory belongs to another pool, or they cannot take cudaMalloc's device-wide
synchronisation).  It keeps its scratch in a fixed array inside the library instead of the per-handle
buffer the regular entry points grow with cudaMalloc.  Checked here through the C ABI against the
regular device entry point and against the CPU oracle, on calls in a row with different swarms (the
shared array keeps the previous call's contents, which must not leak into the next result), including
a swarm large enough that the fixed array forces a coarser split of the draws than the regular path
uses.  Collected last on purpose."""
import numpy as np
import pytest

from tests import workloads as wl

pytestmark = pytest.mark.gpu
RTOL = 1e-10  # the tolerance of the parity tests (BASELINE.json north_star)
MAX_PARTIALS = 262144  # KST_NOALLOC_MAX_PARTIALS (csrc/kst_objective.cu)


def _grid_n_blk(N, S, sm_count, nt=256, minb=2, tile=8):
    """kst_objective_grid (csrc/kst_objective.cu) restated: the draw split the regular path picks."""
    max_chunks = (N + nt - 1) // nt
    cap = sm_count * minb
    tp = tile
    while tp > 1 and ((S + tp - 1) // tp) * max_chunks < cap:
        tp >>= 1
    tiles = (S + tp - 1) // tp
    lo = max(1, (8 * cap + tiles - 1) // tiles)
    best, best_waste = lo, 1e30
    for nb in range(lo, lo + cap):
        tot = nb * tiles
        waves = (tot + cap - 1) // cap
        waste = (waves * cap - tot) / tot
        if waste < best_waste - 1e-12:
            best_waste, best = waste, nb
    nb_dpt = max(N // (nt * 4), max(cap // tiles, 1))
    return max(1, min(best, nb_dpt, max_chunks, 65535))


def test_noalloc_matches_regular_entry_point_and_oracle():
    import torch

    import kirstine_b200 as kb
    import oracle as ko
    assert kb.device_count() >= 1
    dev = torch.device("cuda", 0)
    sm_count = torch.cuda.get_device_properties(dev).multi_processor_count

    def both(h, pts, w):
        S, K = w.shape
        d_pts, d_w = torch.from_numpy(pts).to(dev), torch.from_numpy(w).to(dev)
        d_ref = torch.full((S,), float("nan"), dtype=torch.float64, device=dev)
        d_new = torch.full((S,), float("nan"), dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        h.objective_batch_device(S, K, d_pts.data_ptr(), d_w.data_ptr(), d_ref.data_ptr())
        h.objective_batch_noalloc(S, K, d_pts.data_ptr(), d_w.data_ptr(), d_new.data_ptr())
        h.synchronize()
        return d_ref.cpu().numpy(), d_new.cpu().numpy(), (d_pts, d_w, d_new)

    n = 20_000
    theta = wl.sigemax_prior(n, seed=11)
    h = kb.ProblemHandle(kb._ffi.MODEL_SIGEMAX, [1.0], theta, np.full(n, 1.0 / n), kb._ffi.CRIT_D, None, [0.0], [1.0])
    o = ko.Problem(ko.SIGEMAX, [1.0], theta)
    for call, (S, K, seed) in enumerate([(300, 8, 1), (77, 5, 2)]):
        pts, w = wl.jittered_swarm(S, K, 0.0, 1.0, seed=seed)
        pts[1, 0] = 0.0                               # a dose-0 row, as the swarm's boundary clamp produces
        ref, new, keep = both(h, pts, w)
        assert np.all(np.isfinite(ref)), call
        assert S * _grid_n_blk(n, S, sm_count) <= MAX_PARTIALS
        # same kernels, same draw split, same summation order: the two entry points must agree bit for bit ...
        assert np.array_equal(new, ref), "call %d" % call
        # ... and with the reference arithmetic
        np.testing.assert_allclose(new, o.objective_batch(pts, w, nthreads=8), rtol=RTOL, atol=0, err_msg="call %d" % call)
    # sizes beyond the fixed maximum are refused with a clear error instead of overrunning the array
    with pytest.raises(kb.EngineError, match="exceeds the fixed limit of 4096"):
        h.objective_batch_noalloc(4097, 8, keep[0].data_ptr(), keep[1].data_ptr(), keep[2].data_ptr())

    # a swarm whose regular draw split would need more partial sums than the fixed array holds: the entry point
    # falls back to a coarser split (different summation order, so agreement to rounding, not bit for bit)
    n2, S2, K2 = 80_000, 4000, 5
    assert S2 * _grid_n_blk(n2, S2, sm_count) > MAX_PARTIALS, "this case must take the coarser split"
    theta2 = wl.sigemax_prior(n2, seed=12)
    h2 = kb.ProblemHandle(kb._ffi.MODEL_SIGEMAX, [1.0], theta2, np.full(n2, 1.0 / n2), kb._ffi.CRIT_D, None, [0.0], [1.0])
    pts, w = wl.jittered_swarm(S2, K2, 0.0, 1.0, seed=3)
    for call in range(2):   # twice in a row: the second call finds the first one's partial sums in the array
        if call == 1:       # ... belonging to other particles: the swarm is reversed in between
            pts, w = pts[::-1].copy(), w[::-1].copy()
        ref, new, _ = both(h2, pts, w)
        assert np.all(np.isfinite(ref))
        np.testing.assert_allclose(new, ref, rtol=1e-12, atol=0, err_msg="coarse split, call %d" % call)
    sub = slice(0, S2, 125)   # 32 particles of the last call against the oracle
    want = ko.Problem(ko.SIGEMAX, [1.0], theta2).objective_batch(pts[sub], w[sub], nthreads=8)
    np.testing.assert_allclose(new[sub], want, rtol=RTOL, atol=0)
