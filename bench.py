#!/usr/bin/env python
"""Benchmark of the Bayesian design-objective hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W          # this repo's engine
  python bench.py --impl reference --steps K --warmup W  # CPU restatement of the reference path

Workload = BASELINE.json configs[4] (it fits one GPU): 8-point sigmoid Emax D-design, 10^6 synthetic
prior draws, swarm of 1024 particles.  One "step" = one PSO iteration (swarm move, evaluation of all
particles against all draws, best update) with the swarm resident on the device; the unit of work is
one (particle x prior draw) objective evaluation.  With N GPUs the 10^6 draws are sharded over the
ranks (total work fixed => "strong" scaling) and the S partial objectives are all-reduced per step.

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from tests import workloads as wl  # noqa: E402

METRIC = "Bayesian D objective evals/s (particle x prior draw)"
UNIT = "evals/s"
N_DRAWS, SWARM, K_POINTS = 1_000_000, 1024, 8
FLOPS_PER_UNIT = 441.0   # algorithmic FP64 flops per (particle, draw), config 1b/5
BYTES_PER_DRAW = 8 * (4 + 1)  # compulsory bytes per draw per pass, shared by the tile
WORKLOAD = "cfg5: sigmoid Emax r=4, K=8 on [0,1], DCriterion/Identity, PriorSample 1e6 draws (seed 5), Pso swarm 1024"


def clock_sampler(stop, out, device_index):
    q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active," \
        "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    try:
        p = subprocess.Popen(["nvidia-smi", "-i", str(device_index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                              "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    except OSError:
        return
    def reader():
        for line in p.stdout:
            out.append(line.strip())
    t = threading.Thread(target=reader, daemon=True)
    t.start()
    stop.wait()
    p.terminate()
    t.join(timeout=2)


def summarize_clocks(lines):
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for ln in lines:
        f = [x.strip() for x in ln.split(",")]
        if len(f) < 9:
            continue
        try:
            sm.append(float(f[1]))
            mx.append(float(f[2]))
        except ValueError:
            continue
        for name, val in zip(names, f[5:9]):
            if val.lower().startswith("active"):
                reasons.add(name)
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
    return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_rate(theta, pts, w, sample_particles, sample_draws, threads, repeats=1):
    """Oracle (CPU restatement of the reference path) on a bounded sample of the same workload."""
    import oracle as ko
    sub = theta[:sample_draws]
    prob = ko.Problem(ko.SIGEMAX, [1.0], sub)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        prob.objective_batch(pts[:sample_particles], w[:sample_particles], nthreads=threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return sample_particles * sample_draws / best, best


def run_reference(args):
    """Reference arm: the reference's own CPU algorithm for the path.  The reference is Julia and Julia is
    not installed here or on the GPU box, so this times the C restatement (oracle/, kind "port") with all
    host threads; the stock reference is single-threaded, so this flatters it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    theta = wl.sigemax_prior(N_DRAWS, seed=5)
    pts, w = wl.random_swarm(SWARM, K_POINTS, 0.0, 1.0, seed=5)
    sp, sd = 256, 200_000  # about 2 s per step on 16 host threads
    for _ in range(args.warmup):
        cpu_port_rate(theta, pts, w, sp, sd, threads)
    times = []
    for _ in range(args.steps):
        _, dt = cpu_port_rate(theta, pts, w, sp, sd, threads)
        times.append(dt)
    rate = sp * sd * len(times) / sum(times)
    sample = "%d particles x %d draws of the cfg5 workload per step" % (sp, sd)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------------------------------------------------------
# the other BASELINE.json configs at their own sizes, measured in the same run at N = 1
# ------------------------------------------------------------------------------------------------------------------
FLOPS = {"1": 107.0, "1b": 441.0, "2": 373.0, "3": 221.0, "4": 123.0}   # SURVEY.md 8(d), per unit of work


def _ev_ms(torch, stream, fn, reps, flush=None):
    """Mean device time of fn() over reps calls, CUDA events on the engine's stream, L2 flushed in between."""
    evs = []
    for _ in range(reps):
        if flush is not None:
            with torch.cuda.stream(stream):
                flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        evs.append((e0, e1))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs) / reps


def bench_objective_config(kb, torch, name, desc, model, mp, theta, ub, K, S, iters, peak_tf, flush, crit=0, dt=None,
                           kernel=""):
    n = len(theta)
    h = kb.ProblemHandle(model, mp, theta, np.full(n, 1.0 / n), crit, dt, [0.0], [ub])
    stream = torch.cuda.ExternalStream(h.stream())
    pts, w = wl.random_swarm(S, K, 0.0, ub, seed=5)
    pin_pts, pin_w = torch.from_numpy(pts.copy()).pin_memory(), torch.from_numpy(w.copy()).pin_memory()
    pin_fx = torch.empty(S, dtype=torch.float64).pin_memory()
    e2e_fn = lambda: h.objective_batch_raw(S, K, pin_pts.data_ptr(), pin_w.data_ptr(), pin_fx.data_ptr())  # noqa: E731
    for _ in range(3):
        e2e_fn()
    reps = 20
    h.profile_enable(True)
    t0 = time.perf_counter()
    e2e_ms = _ev_ms(torch, stream, e2e_fn, reps, flush)
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3 / reps
    kern_ms, launches = h.profile_read()
    h.profile_enable(False)
    kern_ms /= max(1, launches)
    # the swarm on the device: `iters` PSO iterations enqueued in one call, timed as a whole
    assert h.pso_init(S, K, pts[0], w[0], seed=5) == 0
    h.pso_run(3)
    pso_ms = _ev_ms(torch, stream, lambda: h.pso_run_async(iters), 3) / iters
    h.synchronize()
    units = float(S) * n
    ach = FLOPS[name] * units / (kern_ms * 1e-3) / 1e12
    out = {"cfg": name, "workload": desc, "unit": UNIT, "units_per_evaluation": units,
           "value": units / (pso_ms * 1e-3), "pso_iteration_us": pso_ms * 1e3, "pso_iterations_timed": 3 * iters,
           "kernel": kernel, "kernel_ms": kern_ms, "kernel_launches_timed": launches,
           "flops_per_unit": FLOPS[name], "achieved_tflops": ach, "frac": ach / peak_tf if peak_tf else None,
           "e2e": {"value": units / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_call": e2e_ms, "wall_ms_per_call": e2e_wall_ms,
                   "h2d_bytes_per_call": int(pin_pts.numel() * 8 + pin_w.numel() * 8), "d2h_bytes_per_call": int(S * 8),
                   "call": "kst_objective_batch (host buffers, pinned)"},
           "l2": "flushed before every timed evaluation call; not inside a pso_run (the iterations of one solve re-read "
                 "the same prior sample, as in the reference)"}
    h.close()
    return out


def bench_gateaux_config(kb, torch, name, desc, theta, G, peak_tf, flush):
    n = len(theta)
    h = kb.ProblemHandle(kb._ffi.MODEL_VU, [0.1], theta, np.full(n, 1.0 / n), 0, None, [0.0], [24.0], m=3,
                         cp_id=kb._ffi.CP_EQUITIME)
    stream = torch.cuda.ExternalStream(h.stream())
    dirs = np.linspace(0.0, 24.0, G)
    t0 = time.perf_counter()
    assert not h.gateaux_prepare([2.0, 8.0, 20.0], [0.3, 0.3, 0.4])
    prep_first_ms = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    for _ in range(5):
        h.gateaux_prepare([2.0, 8.0, 20.0], [0.3, 0.3, 0.4])
    prep_ms = (time.perf_counter() - t0) * 1e3 / 5
    for _ in range(3):
        h.gateaux_eval(dirs)
    reps = 20
    h.profile_enable(True)
    t0 = time.perf_counter()
    e2e_ms = _ev_ms(torch, stream, lambda: h.gateaux_eval(dirs), reps, flush)
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3 / reps
    kern_ms, launches = h.profile_read()
    h.profile_enable(False)
    kern_ms /= max(1, launches)
    units = float(G) * n
    ach = FLOPS["4"] * units / (kern_ms * 1e-3) / 1e12
    out = {"cfg": name, "workload": desc, "unit": "(direction x draw)/s", "units_per_evaluation": units,
           "value": units / (kern_ms * 1e-3), "kernel": "k_gateaux_eval<KstVu,3>", "kernel_ms": kern_ms,
           "kernel_launches_timed": launches, "flops_per_unit": FLOPS["4"], "achieved_tflops": ach,
           "frac": ach / peak_tf if peak_tf else None,
           "gateaux_prepare_ms": prep_ms, "gateaux_prepare_first_call_ms": prep_first_ms,
           "e2e": {"value": units / (e2e_ms * 1e-3), "unit": "(direction x draw)/s", "ms_per_call": e2e_ms,
                   "wall_ms_per_call": e2e_wall_ms, "h2d_bytes_per_call": int(G * 8), "d2h_bytes_per_call": int(G * 8),
                   "call": "kst_gateaux_eval (host buffers)"},
           "l2": "flushed before every timed call"}
    h.close()
    return out


def bench_solves(kb):
    """Whole-solve wall time through the public API (second run of each, after CUDA context and module load)."""
    out = {}
    dp1 = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 10)), model=kb.EmaxModel(1),
                           covariate_parameterization=kb.CopyDose(), prior_knowledge=kb.PriorSample(wl.emax_prior(1000)))
    st1 = kb.DirectMaximization(optimizer=kb.Pso(swarmsize=50, iterations=100),
                                prototype=kb.equidistant_design(kb.region(dp1), 3))
    dp1b = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(dose=(0, 1)), model=kb.SigEmaxModel(1),
                            covariate_parameterization=kb.CopyDose(),
                            prior_knowledge=kb.PriorSample(wl.sigemax_prior(1000)))
    st1b = kb.DirectMaximization(optimizer=kb.Pso(swarmsize=100, iterations=50),
                                 prototype=kb.equidistant_design(kb.region(dp1b), 8))
    dp4 = kb.DesignProblem(criterion=kb.DCriterion(), region=kb.DesignInterval(delta=(0, 24)), model=kb.VUMod(0.1, 3),
                           covariate_parameterization=kb.EquiTime(),
                           prior_knowledge=kb.PriorSample(wl.tpc_prior(1000, seed=4)))
    st4 = kb.Exchange(optimizer_direction=kb.Pso(swarmsize=25, iterations=10),
                      optimizer_weight=kb.Pso(swarmsize=25, iterations=10), steps=8,
                      candidate=kb.uniform_design([[2.0], [8.0], [20.0]]))
    for key, dp, st in (("1", dp1, st1), ("1b", dp1b, st1b), ("4", dp4, st4)):
        kb.seed(1)
        kb.solve(dp, st)
        best = None
        for _ in range(7):   # host wall time of a millisecond-scale call: best of seven
            kb.seed(1)
            t0 = time.perf_counter()
            d, _ = kb.solve(dp, st)
            ms = (time.perf_counter() - t0) * 1e3
            best = ms if best is None else min(best, ms)
        out[key] = {"solve_wall_ms": best, "objective_at_solution": kb.objective(d, dp)}
    return out


def bench_configs(kb, torch, peak_tf, flush):
    res = []
    res.append(bench_objective_config(
        kb, torch, "1", "cfg1: Emax r=3, K=3 on [0,10], D/Identity, 1e3 draws (seed 1), Pso 50 x 100",
        kb._ffi.MODEL_EMAX, [1.0], wl.emax_prior(1000), 10.0, 3, 50, 100, peak_tf, flush,
        kernel="k_objective<KstEmax,D,Identity,uniform>"))
    res.append(bench_objective_config(
        kb, torch, "1b", "cfg1b (tutorial): sigmoid Emax r=4, K=8 on [0,1], D/Identity, 1e3 draws (seed 31415), Pso 100 x 50",
        kb._ffi.MODEL_SIGEMAX, [1.0], wl.sigemax_prior(1000), 1.0, 8, 100, 50, peak_tf, flush,
        kernel="k_objective<KstSigEmax,D,Identity,uniform>"))
    th = wl.sigemax_prior(10_000, seed=2)
    res.append(bench_objective_config(
        kb, torch, "2", "cfg2: sigmoid Emax, K=6, A/DeltaMethod->ED90 (t=1), 1e4 draws (seed 2), Pso 256 x 100",
        kb._ffi.MODEL_SIGEMAX, [1.0], th, 1.0, 6, 256, 100, peak_tf, flush, crit=kb._ffi.CRIT_A,
        dt=wl.ded90_jacobians(th), kernel="k_objective<KstSigEmax,A,DeltaMethod t=1>"))
    res.append(bench_objective_config(
        kb, torch, "3", "cfg3: one-compartment TPC r=3, K=6 on [0,48], D/Identity, 1e5 draws (seed 3), Pso 256 x 100",
        kb._ffi.MODEL_TPC, [1.0], wl.tpc_prior(100_000), 48.0, 6, 256, 100, peak_tf, flush,
        kernel="k_objective<KstTpc,D,Identity,uniform>"))
    res.append(bench_gateaux_config(
        kb, torch, "4", "cfg4: vector unit m=3 (VUMod + EquiTime) on [0,24], D/Identity, Gateaux derivative over a 1e4-point "
        "grid x 1e3 draws (seed 4)", wl.tpc_prior(1000, seed=4), 10_000, peak_tf, flush))
    res.append(bench_gateaux_config(
        kb, torch, "4L", "cfg4 shape at 1e5 draws: the same 1e4-point grid", wl.tpc_prior(100_000, seed=4), 10_000,
        peak_tf, flush))
    solves = bench_solves(kb)
    for r in res:
        if r["cfg"] in solves:
            r["solve"] = dict(solves[r["cfg"]], api="kirstine_b200.solve (" + ("Exchange, 8 steps of Pso(25,10) x 2"
                                                                              if r["cfg"] == "4" else "DirectMaximization") + ")")
    return res


def oracle_parity_check(theta, pts, w, fx, n_check=8):
    """After the timed region: the engine's objective values for a few particles of the e2e swarm against the CPU
    oracle on the FULL prior sample (the oracle as the checker, outside every timed region)."""
    import oracle as ko
    idx = np.unique(np.linspace(0, len(fx) - 1, n_check).astype(int))
    want = ko.Problem(ko.SIGEMAX, [1.0], theta).objective_batch(pts[idx], w[idx], nthreads=os.cpu_count() or 1)
    got = fx[idx]
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin), "engine and oracle disagree on which particles are singular"
    rel = np.abs(got[fin] - want[fin]) / np.abs(want[fin])
    return {"n": int(len(idx)), "n_finite": int(fin.sum()), "max_rel": float(rel.max()) if fin.any() else 0.0,
            "tolerance": 1e-10, "against": "oracle/kirstine_oracle.c on all %d draws, particles %s of the e2e swarm"
                                           % (len(theta), idx.tolist())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--draws", type=int, default=N_DRAWS, help="override the prior size (debugging only)")
    ap.add_argument("--swarm", type=int, default=SWARM)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the `configs` array (the other BASELINE configs)")
    ap.add_argument("--stage", type=int, default=1, choices=[1, 2, 3],
                    help="which launch of the PSO step the engine's own events time (1 = evaluation kernel, the default "
                         "the roofline needs; 2 = move, 3 = best update / multi-GPU tail: for the split of a step in "
                         "profiles/, the roofline keys of such a run are meaningless)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3

    import torch
    import torch.distributed as dist

    import kirstine_b200 as kb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    n_total, S, K = args.draws, args.swarm, K_POINTS
    theta = wl.sigemax_prior(n_total, seed=5)
    j0, j1 = kb.shard_range(n_total, rank, world)
    pts, w = wl.random_swarm(S, K, 0.0, 1.0, seed=5)
    h = kb.ProblemHandle(kb._ffi.MODEL_SIGEMAX, [1.0], theta[j0:j1], np.full(j1 - j0, 1.0 / n_total), kb._ffi.CRIT_D,
                         None, [0.0], [1.0], device=local)
    collective = "none"
    if world > 1:
        # the S partial objectives are summed over NVLink peer memory by one fused kernel per evaluation
        # (kst_comm_peer_*); KST_BENCH_COLLECTIVE=nccl selects the library all-reduce instead
        collective = os.environ.get("KST_BENCH_COLLECTIVE", "peer")
        if collective == "peer":
            ok = 1
            try:
                handles = [None] * world
                dist.all_gather_object(handles, h.comm_peer_handle(world, rank))
                h.comm_peer_open(handles)
            except kb.EngineError as exc:  # no peer access between these GPUs: every rank falls back together
                print("rank %d: peer exchange unavailable (%s), using NCCL" % (rank, exc), file=sys.stderr)
                ok = 0
            flag = torch.tensor([ok], device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 0:
                h.comm_peer_close()
                collective = "nccl"
        if collective == "nccl":
            uid = [kb.ProblemHandle.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            h.comm_init(world, rank, uid[0])
    stream = torch.cuda.ExternalStream(h.stream(), device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(fn, n):
        """n calls of fn, each bracketed by CUDA events on the engine's stream, L2 flushed in between."""
        evs = []
        for _ in range(n):
            with torch.cuda.stream(stream):
                flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn()
            e1.record(stream)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident PSO: `value` ------------------------------------------------------------------
    st = h.pso_init(S, K, pts[0], w[0], seed=5)
    assert st == 0
    for _ in range(args.warmup):
        h.pso_run(1)
    fg_before = h.pso_state()["fg"]
    clock_lines, stop = [], threading.Event()
    sampler = threading.Thread(target=clock_sampler, args=(stop, clock_lines, local), daemon=True)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    h.profile_enable(args.stage)
    barrier()
    ms = timed_steps(lambda: h.pso_run_async(1), args.steps)
    barrier()
    kern_ms, kern_launches = h.profile_read()
    h.profile_enable(False)
    st_after = h.pso_state()  # device -> host read of the step's result (best objective so far)
    assert st_after["fg"] >= fg_before and st_after["n_eval"] == S * (1 + args.warmup + args.steps)
    ms_local = ms
    ms = max_over_ranks(ms)
    units_per_step = float(S) * float(n_total)
    value = units_per_step * args.steps / (ms * 1e-3)

    # ---- end to end through the host-buffer C-ABI call: `e2e` ------------------------------------------
    pin_pts = torch.from_numpy(pts.copy()).pin_memory()
    pin_w = torch.from_numpy(w.copy()).pin_memory()
    pin_fx = torch.empty(S, dtype=torch.float64).pin_memory()
    e2e_fn = lambda: h.objective_batch_raw(S, K, pin_pts.data_ptr(), pin_w.data_ptr(), pin_fx.data_ptr())  # noqa: E731
    for _ in range(args.warmup):
        e2e_fn()
    h.profile_enable(True)
    barrier()
    t0 = time.perf_counter()
    e2e_ms = timed_steps(e2e_fn, args.steps)
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    e2e_kern_ms, e2e_kern_launches = h.profile_read()
    h.profile_enable(False)
    e2e_ms = max_over_ranks(e2e_ms)
    e2e_value = units_per_step * args.steps / (e2e_ms * 1e-3)
    if rank == 0:
        stop.set()
        sampler.join(timeout=3)
    if collective == "peer":
        assert h.comm_peer_error() == 0, "peer exchange timed out"
    fx_check = pin_fx.numpy().copy()
    assert np.all(np.isfinite(fx_check[:1])), "prototype particle must have a finite objective"
    if world > 1:
        # the swarm replicas must have moved in lock-step and every rank must hold the same all-reduced objectives
        st = h.pso_state()
        mine = torch.from_numpy(np.concatenate([st["fx"], st["x_w"].ravel(), [st["fg"]], fx_check])).cuda()
        everyone = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(everyone, mine)
        assert all(torch.equal(everyone[0], t) for t in everyone), "ranks disagree on the swarm state"

    # ---- roofline of the dominant kernel (k_objective) ----------------------------------------------------
    peak_tf = kb.fp64_peak_probe(local)
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    units_per_launch = float(S) * float(j1 - j0)
    kern_avg_ms = kern_ms / max(1, kern_launches)
    achieved_tf = FLOPS_PER_UNIT * units_per_launch / (kern_avg_ms * 1e-3) / 1e12
    alg_bytes = BYTES_PER_DRAW * float(j1 - j0) * (S / 8.0)  # each 8-particle tile streams the shard once
    static_ms = e2e_kern_ms / max(1, e2e_kern_launches)
    static_tf = FLOPS_PER_UNIT * units_per_launch / (static_ms * 1e-3) / 1e12
    traffic, traffic_src = None, None  # DRAM bytes per launch from the committed ncu capture (cannot be measured live)
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        traffic = tj.get("k_objective_cfg5_dram_bytes_per_launch")
        traffic_src = {k: tj.get(k) for k in ("round", "commit", "capture")}
    except OSError:
        pass
    # Headline fraction = the launch in which every particle carries all 8 rows (the e2e leg's static swarm): the flop
    # count charges 8 rows per particle, and in the PSO steps particles clamped to dose 0 legitimately skip that row.
    roofline = {
        "bound": "fp64", "achieved": static_tf, "peak": peak_tf, "unit": "TFLOP/s",
        "frac": static_tf / peak_tf if peak_tf else None, "frac_of_nominal": static_tf / 37.2, "traffic": traffic,
        "traffic_source": traffic_src,
        "kernel": "k_objective<KstSigEmax,D,Identity,uniform>", "kernel_ms": static_ms,
        "launches_timed": e2e_kern_launches, "flops_per_unit": FLOPS_PER_UNIT,
        "measured_on": "static swarm of the e2e leg (all 8 rows live in every particle)",
        "peak_source": "DFMA micro-kernel measured in this run (MEASURED_PEAKS.json has no FP64 figure; "
                       "nominal 148 SM x 64 lanes x 2 x 1.965 GHz = 37.2)",
        "pso_step": {"kernel_ms": kern_avg_ms, "achieved": achieved_tf, "frac": achieved_tf / peak_tf if peak_tf else None,
                     "launches_timed": kern_launches,
                     "note": "same kernel inside the timed PSO steps: particles sitting on dose 0 drop that row, the "
                             "flop count still charges 8 rows, so this figure overstates"},
        "hbm": {"algorithmic_gbs": alg_bytes / (static_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                "note": "served from L2 (shard <= 24 MB); HBM is not the bound"},
    }

    # every rank's own numbers (N > 1): step time between its events and the evaluation kernel inside it.  The step
    # ends with an exchange, so a rank that starts late or runs the kernel slower shows up in everybody's step time.
    per_rank = None
    if world > 1:
        mine = torch.tensor([ms_local / args.steps, kern_avg_ms], dtype=torch.float64, device="cuda")
        everyone = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(everyone, mine)
        per_rank = {"step_ms": [float(t[0]) for t in everyone], "kernel_ms": [float(t[1]) for t in everyone]}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # the N-GPU result against the CPU oracle on the full prior (rank 0, outside every timed region)
    parity = oracle_parity_check(theta, pts, w, fx_check)
    assert parity["max_rel"] <= parity["tolerance"], parity
    configs = None
    if world == 1 and not args.no_configs and args.draws == N_DRAWS and args.swarm == SWARM:
        configs = bench_configs(kb, torch, peak_tf, flush)

    cpu = None
    if world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        sp, sd = 256, 1_000_000  # about 10 s on 16 host threads
        rate, secs = cpu_port_rate(theta, pts, w, sp, sd, threads)
        # the stock reference runs on one thread: the same port on one thread, on 1/16 of the sample
        rate1, secs1 = cpu_port_rate(theta, pts, w, max(1, sp // 16), sd, 1)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "seconds": secs,
               "single_thread": {"value": rate1, "seconds": secs1, "sample": "%d particles x %d draws" % (max(1, sp // 16), sd)},
               "sample": "%d particles x %d draws of the cfg5 workload, oracle/kirstine_oracle.c with OpenMP over "
                         "particles (the Julia reference is single-threaded and not installable here)" % (sp, sd)}

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "draws_total": n_total, "draws_per_gpu": j1 - j0, "swarm": S, "points": K,
                   "step": "one PSO iteration = three launches: k_pso_move, k_objective (all particles x this GPU's draws), and k_pso_best (N = 1) / k_pso_tail (N > 1), which finishes the per-CTA partial sums, sums them over the ranks through NVLink peer memory and does pso_update_best!; with KST_BENCH_COLLECTIVE=nccl: k_pso_move + k_objective + k_obj_finish + ncclAllReduce + k_pso_best",
                   "l2": "flushed between steps (256 MiB write on the same stream, outside the events)",
                   "parallelism": "prior draws sharded over %d GPU(s), swarm replicated" % world,
                   "collective": {"peer": "fused NVLink peer-memory all-reduce kernel (S doubles per step)",
                                  "nccl": "ncclAllReduce (S doubles per step)", "none": "none"}[collective]},
        "clocks": summarize_clocks(clock_lines),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(pin_pts.numel() * 8 + pin_w.numel() * 8),
                "d2h_bytes_per_step": int(S * 8), "ms_per_step": e2e_ms / args.steps,
                "wall_ms_per_step": e2e_wall_ms / args.steps,
                "call": "kst_objective_batch (host buffers, pinned) = one swarm evaluation"},
        # own kernels per timed step on each rank: move, objective, best (N = 1) or tail (N > 1: finish + peer
        # all-reduce + best); on the NCCL path finish and best are separate launches (ncclAllReduce is a library
        # kernel, not counted)
        "gpu_launches": int((4 if collective == "nccl" else 3) * args.steps),
        "roofline": roofline,
        "per_rank": per_rank,
        "stage_timed": {1: "evaluation kernel", 2: "k_pso_move", 3: "k_pso_best / k_pso_tail"}[args.stage],
        "cpu_baseline": cpu,
        "parity_check": parity,
        "configs": configs,
    }
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
