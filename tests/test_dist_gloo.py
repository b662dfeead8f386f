"""N > 1 path on CPU: two gloo ranks shard the prior sample exactly like bench.py / the engine do
(kirstine_b200.shard_range) and all-reduce the S partial objectives.  The oracle stands in for the
per-rank evaluator here (no GPU in this container); what is under test is the partition + collective
logic: shards are disjoint and exhaustive, partial sums add up to the single-rank objective, -Inf and
NaN propagate through the reduction, and every rank ends with identical bits (replicas stay in lock-step)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import kirstine_b200 as kb
    import oracle as ko
    from tests import workloads as wl
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 1001                                        # not divisible by the world size
    theta = wl.sigemax_prior(n, seed=5)
    pts, w = wl.jittered_swarm(12, 8, 0.0, 1.0, seed=5)
    pts[3, :] = 0.5                                 # one distinct point: singular -> -Inf partials on every shard
    j0, j1 = kb.shard_range(n, rank, world)
    prob = ko.Problem(ko.SIGEMAX, [1.0], theta[j0:j1], np.full(j1 - j0, 1.0 / n))
    partial = torch.from_numpy(prob.objective_batch(pts, w))
    dist.all_reduce(partial, op=dist.ReduceOp.SUM)
    np.save(os.path.join(out_dir, "fx_%d.npy" % rank), partial.numpy())
    np.save(os.path.join(out_dir, "range_%d.npy" % rank), np.array([j0, j1]))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_objective_matches_single_rank(tmp_path, world):
    import oracle as ko
    from tests import workloads as wl
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    fx = [np.load(tmp_path / ("fx_%d.npy" % r)) for r in range(world)]
    ranges = [np.load(tmp_path / ("range_%d.npy" % r)) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == 1001
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
    for r in range(1, world):
        assert np.array_equal(fx[0], fx[r], equal_nan=True)            # identical bits on every rank
    theta = wl.sigemax_prior(1001, seed=5)
    pts, w = wl.jittered_swarm(12, 8, 0.0, 1.0, seed=5)
    pts[3, :] = 0.5
    full = ko.Problem(ko.SIGEMAX, [1.0], theta).objective_batch(pts, w)
    assert fx[0][3] == -np.inf and full[3] == -np.inf
    fin = np.isfinite(full)
    np.testing.assert_allclose(fx[0][fin], full[fin], rtol=1e-13)
