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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--draws", type=int, default=N_DRAWS, help="override the prior size (debugging only)")
    ap.add_argument("--swarm", type=int, default=SWARM)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
    h.profile_enable(True)
    barrier()
    ms = timed_steps(lambda: h.pso_run_async(1), args.steps)
    barrier()
    kern_ms, kern_launches = h.profile_read()
    h.profile_enable(False)
    st_after = h.pso_state()  # device -> host read of the step's result (best objective so far)
    assert st_after["fg"] >= fg_before and st_after["n_eval"] == S * (1 + args.warmup + args.steps)
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
    traffic = None  # DRAM bytes per launch from the committed ncu capture of this kernel (cannot be measured live)
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("k_objective_cfg5_dram_bytes_per_launch")
    except OSError:
        pass
    roofline = {
        "bound": "fp64", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
        "frac": achieved_tf / peak_tf if peak_tf else None, "frac_of_nominal": achieved_tf / 37.2, "traffic": traffic,
        "kernel": "k_objective<KstSigEmax,D,Identity>", "kernel_ms": kern_avg_ms, "launches_timed": kern_launches,
        "flops_per_unit": FLOPS_PER_UNIT,
        "peak_source": "DFMA micro-kernel measured in this run (MEASURED_PEAKS.json has no FP64 figure; "
                       "nominal 148 SM x 64 lanes x 2 x 1.965 GHz = 37.2)",
        "static_swarm": {"kernel_ms": e2e_kern_ms / max(1, e2e_kern_launches),
                         "achieved": FLOPS_PER_UNIT * units_per_launch / (e2e_kern_ms / max(1, e2e_kern_launches) * 1e-3) / 1e12,
                         "note": "same kernel on the e2e leg's fixed initial swarm (every particle keeps all 8 rows; in the "
                                 "PSO steps particles clamped to dose 0 drop that row, which the algorithmic count ignores)"},
        "hbm": {"algorithmic_gbs": alg_bytes / (kern_avg_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                "note": "served from L2 (shard <= 24 MB); HBM is not the bound"},
    }

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

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
                   "step": "one PSO iteration: k_pso_move + k_objective + k_pso_best (which also finishes the per-CTA partial sums; N > 1: k_peer_allreduce with the finish fused in, or k_obj_finish + ncclAllReduce, before k_pso_best)",
                   "l2": "flushed between steps (256 MiB write on the same stream, outside the events)",
                   "parallelism": "prior draws sharded over %d GPU(s), swarm replicated" % world,
                   "collective": {"peer": "fused NVLink peer-memory all-reduce kernel (S doubles per step)",
                                  "nccl": "ncclAllReduce (S doubles per step)", "none": "none"}[collective]},
        "clocks": summarize_clocks(clock_lines),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(pin_pts.numel() * 8 + pin_w.numel() * 8),
                "d2h_bytes_per_step": int(S * 8), "ms_per_step": e2e_ms / args.steps,
                "wall_ms_per_step": e2e_wall_ms / args.steps,
                "call": "kst_objective_batch (host buffers, pinned) = one swarm evaluation"},
        # own kernels per timed step on each rank: move, objective, best; the finish of the per-CTA partial sums rides
        # along in k_pso_best on one GPU and in the peer all-reduce kernel on the NVLink path (one more launch there;
        # on the NCCL path k_obj_finish is its own launch and ncclAllReduce is a library kernel, not counted)
        "gpu_launches": int((3 if world == 1 else 4) * args.steps),
        "roofline": roofline,
        "cpu_baseline": cpu,
    }
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
