"""kst_objective_batch_noalloc: the entry point for callers that must not allocate device memory while
they run (their memory belongs to another pool, or they cannot take cudaMalloc's device-wide
synchronisation).  It keeps its scratch in a fixed array inside the library instead of the per-handle
buffer the regular entry points grow with cudaMalloc.  Checked here through the C ABI against the
regular device entry point, on two calls in a row with different swarms (the shared array keeps the
first call's contents, which must not leak into the second result).  Collected last on purpose."""
import numpy as np
import pytest

from tests import workloads as wl

pytestmark = pytest.mark.gpu
RTOL = 1e-10  # the tolerance of the parity tests (BASELINE.json north_star)


def test_noalloc_matches_regular_entry_point_twice():
    import torch

    import kirstine_b200 as kb
    assert kb.device_count() >= 1
    n = 20_000
    theta = wl.sigemax_prior(n, seed=11)
    h = kb.ProblemHandle(kb._ffi.MODEL_SIGEMAX, [1.0], theta, np.full(n, 1.0 / n), kb._ffi.CRIT_D, None, [0.0], [1.0])
    dev = torch.device("cuda", 0)
    for call, (S, K, seed) in enumerate([(300, 8, 1), (77, 5, 2)]):
        pts, w = wl.jittered_swarm(S, K, 0.0, 1.0, seed=seed)
        pts[1, 0] = 0.0                               # a dose-0 row, as the swarm's boundary clamp produces
        d_pts = torch.from_numpy(pts).to(dev)
        d_w = torch.from_numpy(w).to(dev)
        d_ref = torch.full((S,), float("nan"), dtype=torch.float64, device=dev)
        d_new = torch.full((S,), float("nan"), dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        h.objective_batch_device(S, K, d_pts.data_ptr(), d_w.data_ptr(), d_ref.data_ptr())
        h.objective_batch_noalloc(S, K, d_pts.data_ptr(), d_w.data_ptr(), d_new.data_ptr())
        h.synchronize()
        ref, new = d_ref.cpu().numpy(), d_new.cpu().numpy()
        assert np.all(np.isfinite(ref)), call
        np.testing.assert_allclose(new, ref, rtol=RTOL, atol=0, err_msg="call %d" % call)
    # sizes beyond the fixed maximum are refused with a clear error instead of overrunning the array
    with pytest.raises(kb.EngineError, match="exceeds the fixed limit of 4096"):
        h.objective_batch_noalloc(4097, 8, d_pts.data_ptr(), d_w.data_ptr(), d_new.data_ptr())
