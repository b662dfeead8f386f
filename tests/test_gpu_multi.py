"""Two ranks, one GPU each (skipped when fewer than two GPUs are visible): the prior sample is sharded,
the S partial objectives are all-reduced inside the engine (NCCL on the handle's stream), and both swarm
replicas stay in lock-step with each other and with a single-GPU run."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir, peer=False):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    import kirstine_b200 as kb
    from tests import workloads as wl
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)   # plumbing only: carries the 128-byte id
    n = 20_001
    theta = wl.sigemax_prior(n, seed=5)
    j0, j1 = kb.shard_range(n, rank, world)
    h = kb.ProblemHandle(2, [1.0], theta[j0:j1], np.full(j1 - j0, 1.0 / n), 0, None, [0.0], [1.0], device=rank)
    if peer:   # NVLink peer-memory exchange instead of NCCL
        handles = [None] * world
        dist.all_gather_object(handles, h.comm_peer_handle(world, rank))
        h.comm_peer_open(handles)
    else:
        uid = [kb.ProblemHandle.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        h.comm_init(world, rank, uid[0])
    pts, w = wl.jittered_swarm(40, 8, 0.0, 1.0, seed=5)
    fx = h.objective_batch(pts, w)
    assert h.pso_init(40, 8, pts[0], w[0], seed=11) == 0
    tfx, tp, tw = h.pso_run(6)
    st = h.pso_state()
    h.gateaux_prepare(pts[0], w[0])
    gd = h.gateaux_eval(np.linspace(0, 1, 33))
    # the Exchange strategy's direction swarm (Gateaux mode) and a whole Exchange run on the sharded prior
    _, gfx, gpx, _, _ = h.pso_optimize(12, 1, 5, pts[0][[1, 4, 6]].reshape(3, 1, 1), np.ones((3, 1)), seed=3,
                                       mode=kb._ffi.PSO_MODE_GATEAUX, state0=False)
    ex = h.exchange_run(pts[0], w[0], 2, (12, 4, 2.05, 2.05), (10, 4, 2.05, 2.05), seed=8, state0=False)
    # a swarm of several 128-particle slices: the tail kernel exchanges and updates the bests slice by slice
    assert h.pso_init(300, 8, pts[0], w[0], seed=12) == 0
    btfx, _, _ = h.pso_run(3)
    bst = h.pso_state()
    if peer:
        assert h.comm_peer_error() == 0
    np.savez(os.path.join(out_dir, "r%d.npz" % rank), fx=fx, tfx=tfx, tp=tp, tw=tw, x=st["x_pts"], xw=st["x_w"], gd=gd,
             gfx=gfx, gpx=gpx, exw=ex["weights"], exp=ex["points"], exfx=ex["w_fx"], btfx=btfx, bx=bst["x_pts"],
             bfx=bst["fx"], bg=bst["g_pts"])
    dist.destroy_process_group()


@pytest.mark.parametrize("peer", [False, True], ids=["nccl", "nvlink-peer"])
def test_two_gpu_sharded_run_matches_single_gpu(tmp_path, peer):
    import torch
    import torch.multiprocessing as mp

    import kirstine_b200 as kb
    from tests import workloads as wl
    if kb.device_count() < 2:
        pytest.skip("needs two GPUs")
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path), peer), nprocs=2, join=True)
    a, b = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    for key in a.files:
        assert np.array_equal(a[key], b[key]), key                       # replicas bit-identical
    n = 20_001
    theta = wl.sigemax_prior(n, seed=5)
    h = kb.ProblemHandle(2, [1.0], theta, np.full(n, 1.0 / n), 0, None, [0.0], [1.0])
    pts, w = wl.jittered_swarm(40, 8, 0.0, 1.0, seed=5)
    np.testing.assert_allclose(a["fx"], h.objective_batch(pts, w), rtol=1e-12)
    assert h.pso_init(40, 8, pts[0], w[0], seed=11) == 0
    tfx, tp, tw = h.pso_run(6)
    np.testing.assert_allclose(a["tfx"], tfx, rtol=1e-11)
    np.testing.assert_allclose(a["x"], h.pso_state()["x_pts"], rtol=1e-9, atol=1e-12)
    h.gateaux_prepare(pts[0], w[0])
    np.testing.assert_allclose(a["gd"], h.gateaux_eval(np.linspace(0, 1, 33)), rtol=1e-10, atol=1e-12)
    _, gfx, gpx, _, _ = h.pso_optimize(12, 1, 5, pts[0][[1, 4, 6]].reshape(3, 1, 1), np.ones((3, 1)), seed=3,
                                       mode=kb._ffi.PSO_MODE_GATEAUX, state0=False)
    np.testing.assert_allclose(a["gfx"], gfx, rtol=1e-9, atol=1e-10 * np.abs(gfx).max())
    np.testing.assert_allclose(a["gpx"], gpx, rtol=1e-9)
    ex = h.exchange_run(pts[0], w[0], 2, (12, 4, 2.05, 2.05), (10, 4, 2.05, 2.05), seed=8, state0=False)
    np.testing.assert_allclose(a["exfx"], ex["w_fx"], rtol=1e-10)
    np.testing.assert_allclose(a["exp"], ex["points"], rtol=1e-8)
    np.testing.assert_allclose(a["exw"], ex["weights"], rtol=1e-7, atol=1e-9)
    assert h.pso_init(300, 8, pts[0], w[0], seed=12) == 0
    btfx, _, _ = h.pso_run(3)
    bst = h.pso_state()
    np.testing.assert_allclose(a["btfx"], btfx, rtol=1e-11)
    np.testing.assert_allclose(a["bfx"], bst["fx"], rtol=1e-9)
    np.testing.assert_allclose(a["bx"], bst["x_pts"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(a["bg"], bst["g_pts"], rtol=1e-9, atol=1e-12)
    # the sharded result against the CPU oracle on the whole prior
    import oracle as ko
    want = ko.Problem(ko.SIGEMAX, [1.0], theta).objective_batch(pts, w, nthreads=8)
    np.testing.assert_allclose(a["fx"], want, rtol=1e-10)
