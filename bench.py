#!/usr/bin/env python
"""bench.py — batched LQ solves/sec (ProxRiccati backward + forward) on B200.

Contract (one JSON line on stdout, printed by rank 0):
  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c1b|c3|c4|c5] [--impl ours|reference]
  N > 1: launched by torchrun, one rank per GPU; the batch dimension is sharded (no
  collective inside the sweep); one NCCL gather of the trajectories per step (north_star).

A "step" = one backward + forward over the whole per-GPU batch of synthetic problems.
  value     solves/s with the problem already resident in HBM (device generator, seed 42)
  e2e       solves/s through the C ABI with HOST buffers: H2D of the packed problem, sweep,
            D2H of (xs, us, vs, lbdas), all inside the timed region
  roofline  dominant kernel (backward sweep): algorithmic bytes / CUDA-event duration vs the
            measured HBM peak (small-nx, HBM-bound) — flops vs measured FP64 peak also printed
  cpu_baseline  the CPU oracle port (oracle/, OpenMP over problems) on a bounded sample
`--impl reference` times that CPU port alone (the reference itself needs Eigen: not buildable
here) on this repo's config/metric.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

MU = 1e-11  # bench/gar-riccati.cpp:26

# BASELINE.json configs. batch = per-GPU batch (weak scaling: every rank gets the full config batch).
WORKLOADS = {
    "c2": dict(nx=12, nu=4, nc=0, nth=0, T=50, batch=16384, legs=0,
               desc="configs[1]: 16384 independent random LQ problems nx=12 nu=4 T=50"),
    "c1b": dict(nx=36, nu=12, nc=0, nth=0, T=100, batch=4096, legs=0,
                desc="north_star target shape nx=36 nu=12 T=100, batch 4096"),
    "c1": dict(nx=36, nu=12, nc=0, nth=0, T=100, batch=1, legs=0,
               desc="configs[0]: single problem nx=36 nu=12 T=100 (latency)"),
    "c3": dict(nx=36, nu=12, nc=6, nth=4, T=200, batch=4096, legs=0,
               desc="configs[2]: nc=6 nth=4 nx=36 nu=12 T=200 batch 4096"),
    "c4": dict(nx=36, nu=12, nc=0, nth=0, T=4096, batch=1, legs=64,
               desc="configs[3]: parallel-in-time T=4096, 64 legs, single problem"),
    "c5": dict(nx=76, nu=32, nc=0, nth=0, T=100, batch=1024, legs=0,
               desc="configs[4]: Talos-shaped nx=76 nu=32 T=100 batch 1024"),
    "p30": dict(nx=30, nu=10, nc=0, nth=0, T=100, batch=4096, legs=0,
                desc="off-grid shape nx=30 nu=10 T=100 batch 4096: no instantiated sweep, embedded in (36, 12) by exact padding"),
}


def flops_model(n, m, c, h, T, c0):
    """dense op count of the reference algorithm per solve."""
    fs = (17 * n**3 + 10 * n * n * m + 2 * m * m * n + (m + c) ** 3 / 3 + 2 * (m + c) ** 2 * (n + 1)
          + 2 * n * n * c + 10 * n * n + 6 * n * m + 2 * n * c)
    if h > 0:
        fs += 8 * n * n * h + 8 * n * m * h + 2 * (m + c) ** 2 * h + 2 * h * h * (n + m) + 2 * h * (n + m)
    ff = 2 * (m + c + 2 * n) * (n + h)
    fi = (n + c0) ** 3 / 3 + 2 * (n + c0) ** 2 * (1 + h)
    return T * (fs + ff) + fi, T * fs + fi, T * ff


def bytes_model(n, m, c, h, T):
    """algorithmic bytes per knot: in / gains / out (doubles * 8)."""
    inn = 3 * n * n + 2 * n * m + m * m + 2 * n + m + c * (n + m + 1)
    if h > 0:
        inn += h * h + n * h + m * h + h
    gains = (m + c + 2 * n) * (n + 1 + h)
    out = 2 * n + m + c
    return 8 * inn * (T + 1), 8 * gains * (T + 1), 8 * out * (T + 1)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.rows, self._halt = device, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.splitlines()[0].split(",")])
            except Exception:
                pass
            self._halt.wait(0.1)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def cpu_port_throughput(w, budget_s=12.0, dist="ref"):
    """Time the CPU oracle port (OpenMP over independent problems, all host cores) on a bounded
    sample of the workload. -> (solves/s, cores, sample description)."""
    import garb200_pkg
    garb200_pkg.load()
    from aligator_swk_b200.synth import generate_problems
    import pyoracle as po
    po.build()
    # all host cores this process may use — NOT OMP_NUM_THREADS, which torchrun pins to 1
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    nx, nu, nc, nth, T, legs = w["nx"], w["nu"], w["nc"], w["nth"], w["T"], w["legs"]
    rng = np.random.default_rng(3)

    def run(nprob, reps):
        pp = generate_problems(nprob, T, nx, nu, nc, nth, seed=42, dist=dist)
        theta = rng.standard_normal(nth) if nth else None
        secs, used = po.bench(pp, MU, MU, theta=theta, num_legs=legs, reps=reps, nthreads=cores)
        return secs, used

    if legs >= 2 or w["batch"] == 1:
        secs, used = run(1, 1)                      # warm-up
        reps = int(max(3, min(200, budget_s / max(secs, 1e-6))))
        secs, used = run(1, reps)
        return reps / secs, used, f"1 problem x {reps} repetitions, {used} threads" + (f" ({legs} legs)" if legs >= 2 else "")
    n0 = max(cores, 8)
    secs, used = run(n0, 1)
    per = secs / n0 * used                          # core-seconds per solve
    nprob = int(max(n0, min(w["batch"], 4096, budget_s * used / per / 2)))
    reps = 2
    secs, used = run(nprob, reps)
    return nprob * reps / secs, used, f"{nprob} of {w['batch']} problems x {reps} repetitions, OpenMP over problems, {used} threads"


def reference_arm(args, w, rank, world):
    """--impl reference: the reference's CPU path.  The real reference needs Eigen 3.4 (absent,
    no network): the Eigen-free oracle port is timed instead (kind = "port").
    One bounded sample of the workload is generated ONCE; every step is one timed pass over it with all host
    threads, sized so that the whole --steps K --warmup W run stays within ~2 minutes whatever K is."""
    if rank != 0:
        return
    import garb200_pkg
    garb200_pkg.load()
    from aligator_swk_b200.synth import generate_problems
    import pyoracle as po
    po.build()
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    nx, nu, nc, nth, T, legs = w["nx"], w["nu"], w["nc"], w["nth"], w["T"], w["legs"]
    theta = np.random.default_rng(3).standard_normal(nth) if nth else None
    single = legs >= 2 or w["batch"] == 1
    nsteps = max(1, args.steps + args.warmup)
    per_step_budget = max(0.25, min(15.0, 90.0 / nsteps))          # seconds of CPU work per step
    n0 = 1 if single else max(cores, 8)
    pp = generate_problems(n0, T, nx, nu, nc, nth, seed=42, dist=args.dist)
    secs, used = po.bench(pp, MU, MU, theta=theta, num_legs=legs, reps=1, nthreads=cores)   # calibration (+ warm caches)
    if single:
        nprob, reps = 1, int(max(1, min(500, per_step_budget / max(secs, 1e-6))))
    else:
        per = secs / n0                                             # wall seconds per solve at full thread count
        nprob = int(max(n0, min(w["batch"], 1024, per_step_budget / max(per, 1e-9))))
        reps = int(max(1, min(50, per_step_budget / max(per * nprob, 1e-9))))
        if nprob != n0:
            pp = generate_problems(nprob, T, nx, nu, nc, nth, seed=42, dist=args.dist)
    per_step = []
    for i in range(args.warmup + args.steps):
        secs, used = po.bench(pp, MU, MU, theta=theta, num_legs=legs, reps=reps, nthreads=cores)
        if i >= args.warmup:
            per_step.append(nprob * reps / secs)
    cores = used
    sample = (f"{nprob} of {w['batch']} problems x {reps} repetition(s) per step, generated once; OpenMP over problems, {used} threads"
              if not single else f"1 problem x {reps} repetitions per step, {used} threads" + (f" ({legs} legs)" if legs >= 2 else ""))
    v = float(np.mean(per_step))
    ms = 1e3 * w["batch"] * world / v
    line = {
        "impl": "reference", "metric": "batched LQ solves/sec (bwd+fwd)", "value": v, "unit": "solves/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args.workload, w, world, args.dist),
        "knots_per_s": v * (w["T"] + 1),
        "cpu_baseline": {"value": v, "unit": "solves/s", "cores": cores, "kind": "port", "sample": sample,
                         "gflops_per_core": v * flops_model(w["nx"], w["nu"], w["nc"], w["nx"] if w["legs"] >= 2 else w["nth"], w["T"], w["nx"])[0] / max(cores, 1) / 1e9},
        "e2e": {"value": v, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


DIST_DESC = {"ref": "ref (tests/gar/test_util.cpp distributions, distinct knots)",
             "explicit": "explicit-dyn (E = -I, A = 0.9 I + 0.1/sqrt(n) N(0,1), rest as ref; SURVEY 8(d) secondary row). NOTE: the ref generator's zeroed Q block makes the stage costs nonconvex; with E = -I that propagates into Rhat, so every knot takes the Bunch-Kaufman fallback",
             "explicit_psd": "explicit-dyn with convex stage costs (E = -I, A = 0.9 I + 0.1/sqrt(n) N(0,1), [Q S; S^T R] = W W^T without the zeroed block): the shape of what ProxDDP feeds (explicit-dynamics.hxx:32, PSD Gauss-Newton Hessians)"}
DIST_ID = {"ref": 0, "explicit": 1, "explicit_psd": 2}


def workload_config(name, w, world, dist="ref"):
    return {"workload": f"{name}: {w['desc']}" + (f" [{dist}: E = -I]" if dist != "ref" else ""), "nx": w["nx"], "nu": w["nu"], "nc": w["nc"], "nth": w["nth"],
            "horizon": w["T"], "batch_per_gpu": w["batch"], "global_batch": w["batch"] * world,
            "num_legs": w["legs"], "mu": MU, "dist": DIST_DESC[dist],
            "l2_policy": "inputs larger than L2 (126 MB) for the batched workloads; none needed for latency rows",
            "parallelism": f"batch-sharded x{world}"}


def parity_sample(s, w, theta_np, kkt, k=8):
    """Measured parity at bench scale: the k problems of THIS rank's timed batch with
    the worst KKT residual are downloaded and re-solved by the CPU oracle on the very same bytes.
    -> dict(max_rel_err over (xs, us, vs, lbdas), kkt of the GPU and of the oracle on those problems)."""
    from aligator_swk_b200.synth import PackedProblems
    from aligator_swk_b200.layout import KnotLayout
    import pyoracle as po
    po.build()
    nx, nu, nc, nth, T, legs = w["nx"], w["nu"], w["nc"], w["nth"], w["T"], w["legs"]
    xs, us, vs, lb = s.forward(theta_np)
    worst = np.argsort(-np.nan_to_num(kkt.max(axis=1), nan=np.inf))[:k]
    L = KnotLayout(nx, nu, nc, nth)
    err, kg, ko = 0.0, 0.0, 0.0
    for b in worst:
        b = int(b)
        knots, G0, g0 = s.download_packed(b, 1)
        pp = PackedProblems(L, T, s.nc0, s.nu_t, s.nc_t, knots, G0, g0)
        th = None if theta_np is None else theta_np[b]
        # the oracle's SERIAL solver is the yardstick also for the parallel-in-time workload (tests/gar/parallel.cpp:243-260)
        ref = po.prox_solve(pp, MU, MU, theta=th)
        for got, key in ((xs[b], "xs"), (us[b], "us"), (vs[b], "vs"), (lb[b], "lbdas")):
            if got.size:
                err = max(err, float(np.abs(got - ref[key]).max() / max(np.abs(ref[key]).max(), 1e-300)))
        kg = max(kg, float(kkt[b].max())); ko = max(ko, float(ref["kkt"].max()))
    out = {"problems": int(len(worst)), "selection": "worst KKT residual of this rank's timed batch",
           "max_rel_err": err, "kkt_gpu": kg, "kkt_oracle": ko, "tolerance": 1e-10}
    # Which one moved?  For the single worst problem: GPU and oracle against the dense KKT system of the
    # same bytes solved with extended-precision iterative refinement (oracle/extended_precision.py).
    rows = s.nc0 + (T + 1) * nx + T * (nu + nc) + T * nx
    if nth == 0 and rows <= 12000 and len(worst):
        try:
            from extended_precision import refined_solve
            b = int(worst[0])
            knots, G0, g0 = s.download_packed(b, 1)
            pp = PackedProblems(L, T, s.nc0, s.nu_t, s.nc_t, knots, G0, g0)
            mat, rhs = po.dense_kkt(pp, MU, MU)
            z, hist = refined_solve(mat, rhs)
            rx, ru, rv, rl = po.dense_solution_to_traj(pp, z)
            ref = po.prox_solve(pp, MU, MU)
            rel = lambda a, r_: float(np.abs(a - r_).max() / max(np.abs(r_).max(), 1e-300)) if a.size else 0.0
            out["worst_problem_vs_extended_precision"] = {
                "gpu": max(rel(xs[b], rx), rel(us[b], ru), rel(vs[b], rv), rel(lb[b], rl)),
                "oracle": max(rel(ref["xs"], rx), rel(ref["us"], ru), rel(ref["vs"], rv), rel(ref["lbdas"], rl)),
                "refinement_backward_error": float(min(hist)), "kkt_rows": int(rows)}
        except Exception as e:   # evidence row only
            out["worst_problem_vs_extended_precision"] = {"error": str(e)[:200]}
    return out


def gpu_arm(args, w, rank, world, local_rank):
    import torch
    import garb200_pkg
    garb200_pkg.load()
    from aligator_swk_b200 import capi, sharding

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    nx, nu, nc, nth, T, B, legs = w["nx"], w["nu"], w["nc"], w["nth"], w["T"], w["batch"], w["legs"]
    if args.scaling == "strong":          # the config's batch is the GLOBAL batch: contiguous slices
        lo, hi = sharding.shard_bounds(B, rank, world)
        B = hi - lo
        w = dict(w, batch=B)
    K, W = args.steps, args.warmup
    s = capi.Solver(nx, nu, nc, nth, horizon=T, batch=B, num_legs=legs, device=local_rank)
    s.set_tuning(**args.tune_knobs)
    s.generate(seed=42 + rank, dist=DIST_ID[args.dist])
    s.synchronize_stream_only()
    theta_d = None
    if nth:
        theta_t = torch.randn(B, nth, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
        theta_d = theta_t.data_ptr()
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # trajectory gather to rank 0 (north_star: a single NCCL gather, no collective inside the
    # sweep): the forward kernel writes straight into the send buffer.
    send = recv = None
    out_ptrs = [{}, {}]
    p2p_dst = None
    gather_kind = None
    if dist is not None:
        sizes = sharding.trajectory_sizes(B, T, nx, nu, nc)
        # two send buffers: the forward rollout of step i+1 writes the other one while step i's is in flight
        nmax = torch.tensor([sum(sizes)], dtype=torch.int64, device="cuda")
        dist.all_reduce(nmax, op=dist.ReduceOp.MAX)
        slot = int(nmax.item())            # strong scaling: slices differ by at most one problem
        send = [torch.zeros(slot, dtype=torch.float64, device="cuda") for _ in range(2)]
        recv_flat = torch.empty(world * slot, dtype=torch.float64, device="cuda") if rank == 0 else None
        recv = [recv_flat[r * slot: r * slot + slot] for r in range(world)] if rank == 0 else None
        gather_kind = "nccl"
        if args.gather in ("auto", "p2p"):
            # SM-free gather: rank 0 exports its receive buffer (CUDA IPC), every rank maps it and copies its
            # slice there with a plain device-to-device copy (copy engines over NVLink) on a side stream
            ok, handle = 1, [None]
            try:
                if rank == 0:
                    handle = [recv_flat.untyped_storage()._share_cuda_()]
            except Exception:
                ok = 0
            dist.broadcast_object_list(handle, src=0)
            try:
                if rank != 0 and handle[0] is not None:
                    st = torch.UntypedStorage._new_shared_cuda(*handle[0])
                    full = torch.empty(0, dtype=torch.float64, device=st.device).set_(st)
                    p2p_dst = full[rank * slot: rank * slot + slot]
                    p2p_dst.copy_(send[0])          # exercises the peer mapping once
                    torch.cuda.synchronize()
                elif rank == 0:
                    p2p_dst = recv_flat[:slot]
                else:
                    ok = 0
            except Exception:
                ok = 0
            flag = torch.tensor([ok], dtype=torch.int32, device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 1:
                gather_kind = "p2p"
            else:
                p2p_dst = None
                if args.gather == "p2p":
                    raise SystemExit("bench.py: --gather p2p requested but the CUDA IPC mapping failed on some rank")
        for k in range(2):
            for name, view in zip(("d_xs", "d_us", "d_vs", "d_lbdas"), sharding.carve(send[k], sizes)):
                out_ptrs[k][name] = view.data_ptr()

    # The gather of step i runs on a side stream and overlaps the backward sweep AND the forward rollout
    # of step i+1 (double-buffered send buffer); a forward only waits for the gather that last read its buffer.
    side = torch.cuda.Stream() if dist is not None else None
    gathered = [None, None]
    stepno = [0]

    def step_device():
        k = stepno[0] & 1
        s.backward(MU, MU)
        if side is not None and gathered[k] is not None:
            torch.cuda.current_stream().wait_event(gathered[k])
        s.forward_device(d_theta=theta_d, **out_ptrs[k])

    def gather():
        k = stepno[0] & 1
        stepno[0] += 1
        if dist is not None:
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                if p2p_dst is not None:
                    p2p_dst.copy_(send[k], non_blocking=True)
                else:
                    sharding.gather_to_root(dist, send[k], rank, world, recv)
                ev = torch.cuda.Event()
                ev.record(side)
                gathered[k] = ev

    def drain():
        if side is not None:
            torch.cuda.current_stream().wait_stream(side)

    # ---- value: device-resident inputs -> device-resident outputs -----------------------
    for _ in range(W):
        step_device(); gather()
    drain()
    barrier()
    l0 = s.kernel_launches()
    s.timing_log_start(K)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step_device(); gather()
    drain()                              # the last gather is inside the timed region
    e1.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    ms_total = e0.elapsed_time(e1)
    launches = s.kernel_launches() - l0
    bwd_ms, fwd_ms = s.timing_log_read()
    s.forward_device(d_theta=theta_d)    # into the solver-owned trajectory (scored below)
    s.synchronize()                      # raises if any problem flagged a numerical failure
    theta_np = None if not nth else theta_t.cpu().numpy()
    kkt = s.kkt_error(MU, MU, theta=theta_np)
    par = parity_sample(s, w, theta_np, kkt, k=args.parity_k) if args.parity_k > 0 else None
    # every rank scores ITS slice: maxima over all ranks (round 1 reported rank 0's only)
    red = torch.tensor([ms_total, float(np.nanmax(kkt[:, 0])), float(np.nanmax(kkt[:, 1])), float(np.nanmax(kkt[:, 2]))]
                       + ([par["max_rel_err"], par["kkt_gpu"], par["kkt_oracle"]] if par else [0.0, 0.0, 0.0]),
                       dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
    red = red.cpu().numpy()
    ms_total = float(red[0])
    kkt_max = {"dyn": float(red[1]), "cstr": float(red[2]), "dual": float(red[3]), "over": f"all {world} rank(s)"}
    if par:
        par.update(max_rel_err=float(red[4]), kkt_gpu=float(red[5]), kkt_oracle=float(red[6]), ranks=world,
                   problems=par["problems"] * world)
    ms_step = ms_total / K
    btot = torch.tensor([B], dtype=torch.int64, device="cuda")
    if dist is not None:
        dist.all_reduce(btot)
    value = int(btot.item()) / (ms_step * 1e-3)

    # ---- e2e: host buffers through the C ABI, H2D + D2H inside the timed region ----------
    e2e = None
    if not args.no_e2e:
        e2e = e2e_run(torch, capi, s, w, max(3, min(K, 10)), W, dist, world, theta_np)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (backward sweep) --------------------------------
    c0 = nx
    f_all, f_bwd, f_fwd = flops_model(nx, nu, nc, nx if legs >= 2 else nth, T, c0)
    b_in, b_gain, b_out = bytes_model(nx, nu, nc, nx if legs >= 2 else nth, T)
    bwd = float(np.mean(bwd_ms)) * 1e-3
    fwd = float(np.mean(fwd_ms)) * 1e-3
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    fp64_dfma, fp64_dmma = capi.fp64_peak(local_rank, 0), capi.fp64_peak(local_rank, 1)
    fp64_peak = max(fp64_dfma, fp64_dmma)
    ach_gbs = B * (b_in + b_gain) / bwd / 1e9
    ach_tf = B * f_bwd / bwd / 1e12
    hbm_bound = (ach_gbs / hbm_peak) >= (ach_tf / fp64_peak)
    roof = {"kernel": "backward pass = E^-1 pre-pass + sweep kernel (2 launches per step; generic shapes: 1)", "bound": "hbm" if hbm_bound else "tensor",
            "bound_detail": "HBM bytes" if hbm_bound else "FP64 on the tensor pipe (DMMA m8n8k4; no tcgen05 kind for f64)",
            "achieved": ach_gbs if hbm_bound else ach_tf, "peak": hbm_peak if hbm_bound else fp64_peak,
            "unit": "GB/s" if hbm_bound else "TFLOP/s",
            "frac": (ach_gbs / hbm_peak) if hbm_bound else (ach_tf / fp64_peak), "traffic": None,
            "peak_source": ("MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s") if hbm_bound
            else "measured in this run: max(DFMA loop, DMMA m8n8k4 loop)",
            "hbm_gbs": ach_gbs, "hbm_frac": ach_gbs / hbm_peak, "fp64_tflops": ach_tf, "fp64_frac": ach_tf / fp64_peak,
            "fp64_peak_dfma": fp64_dfma, "fp64_peak_dmma": fp64_dmma,
            "kernel_ms": bwd * 1e3, "forward_kernel_ms": fwd * 1e3,
            "forward_gbs": B * (b_gain + b_out) / fwd / 1e9,
            "algorithmic_bytes_per_launch": B * (b_in + b_gain), "algorithmic_flops_per_launch": B * f_bwd,
            "kernel_share_of_step": (bwd * 1e3) / ms_step}
    # DRAM traffic of the backward pass from the committed `ncu --set full` capture of this same
    # workload at this same batch (profiles/rNN_ncu_digest.json, written by tools/make_profiles.py)
    try:
        import glob
        digests = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_digest.json")))
        full = (json.load(open(digests[-1])).get(args.workload) or {}).get("full") or {}
        parts = {k: v["dram_bytes"] for k, v in full.items() if k.startswith("einv") or k.startswith("sweep")}
        if parts and any(k.startswith("sweep") for k in parts) and B == WORKLOADS[args.workload]["batch"]:
            roof["traffic"] = float(sum(parts.values()))
            roof["traffic_source"] = os.path.basename(digests[-1]) + ": dram__bytes_read.sum + dram__bytes_write.sum of " + " + ".join(sorted(parts))
    except Exception:
        pass
    cpu_v, cores, sample = cpu_port_throughput(w, budget_s=args.cpu_budget, dist=args.dist)
    line = {
        "metric": "batched LQ solves/sec (bwd+fwd)", "value": value, "unit": "solves/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.workload, w, world, args.dist),
        "explicit_path": bool(s.used_explicit_path()),
        "knots_per_s": value * (T + 1),
        "kkt_max": kkt_max, "parity_sample": par,
        "roofline": roof,
        "cpu_baseline": {"value": cpu_v, "unit": "solves/s", "cores": cores, "kind": "port", "sample": sample,
                         "gflops_per_core": cpu_v * f_all / max(cores, 1) / 1e9,
                         "note": "Eigen-free port of the reference algorithm (oracle/): malloc-free stage solve, register-blocked GEMM, 4-RHS blocked triangular solves; the reference itself needs Eigen 3.4 (not installable here)"},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        "collective": None if world == 1 else (
            "one gather of (xs,us,vs,lbdas) to rank 0 per step on a side stream (overlaps the next backward sweep): "
            + ("device-to-device copies into rank 0's buffer through a CUDA IPC mapping — copy engines over NVLink, no kernel on any SM"
               if gather_kind == "p2p" else "NCCL gather")),
        "gather": gather_kind,
    }
    if world == 1 and not args.no_e2e:
        line["consumers"] = consumer_bandwidth(torch, s, w, hbm_peak)
        line["e2e_dropin"] = dropin_latency()
    if world == 1 and args.others:
        s.close()
        line["other_configs"] = {}
        for spec in args.others.split(","):
            name, _, dname = spec.partition("+")
            dname = dname or "ref"
            if name in WORKLOADS and dname in DIST_DESC and (name, dname) != (args.workload, args.dist):
                line["other_configs"][spec] = short_run(torch, capi, name, local_rank, hbm_peak, fp64_peak, dist=dname)
    line["config"]["global_batch"] = int(btot.item())
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def dropin_latency():
    """c1 (1 x nx=36 nu=12 T=100) end to end through the ADAPTER's call sequence (garb200_stage_knot x (T+1),
    commit, backward, forward, updateGains read-back), timed by tools/dropin_bench (built by build())."""
    exe = os.path.join(ROOT, "tools", "dropin_bench")
    if not os.path.exists(exe):
        return {"error": "tools/dropin_bench not built (run __graft_entry__.build())"}
    try:
        out = subprocess.run([exe, "36", "12", "100", "50"], capture_output=True, text=True, timeout=120)
        return json.loads(out.stdout.strip().splitlines()[-1])
    except Exception as e:
        return {"error": str(e)[:200]}


def consumer_bandwidth(torch, s, w, hbm_peak, K=5):
    """The device-side steps around the solve on the bench batch: HBM-bound kernels,
    algorithmic bytes / CUDA-event time."""
    nx, nu, nc, T, B, legs = w["nx"], w["nu"], w["nc"], w["T"], w["batch"], w["legs"]
    if legs >= 2:
        return None
    try:
        n_x, n_u, n_v, n_l = B * (T + 1) * nx, B * (T + 1) * nu, B * (T + 1) * nc, B * (nx + T * nx)
        mk = lambda n: torch.zeros(max(n, 1), dtype=torch.float64, device="cuda")
        cur = [mk(n) for n in (n_x, n_u, n_v, n_l)]
        out = [mk(n) for n in (n_x, n_u, n_v, n_l)]
        ptr = lambda ts, ns: [t.data_ptr() if n else None for t, n in zip(ts, ns)]
        ns = (n_x, n_u, n_v, n_l)
        res = {}
        def timed(fn):
            fn(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(K):
                fn()
            e1.record(); torch.cuda.synchronize()
            return e0.elapsed_time(e1) / K * 1e-3
        t = timed(lambda: s.linear_step(0.5, ptr(cur, ns), ptr(out, ns)))
        by = 3 * 8 * sum(ns)
        res["linear_step"] = {"ms": t * 1e3, "gbs": by / t / 1e9, "hbm_frac": by / t / 1e9 / hbm_peak}
        t = timed(lambda: s.rollout_device(0.5, cur[0].data_ptr(), out[1].data_ptr() if nu else None,
                                           out[2].data_ptr() if nc else None, out[3].data_ptr()))
        rows = nu + nc + nx
        by = 8 * B * (T + 1) * (rows * (nx + 1) + nx + rows)
        res["rollout_device"] = {"ms": t * 1e3, "gbs": by / t / 1e9, "hbm_frac": by / t / 1e9 / hbm_peak,
                                 "bytes_model": "ff + K/Z/L rows of fb read, dxs read, dus/dvs/dlams written"}
        return res
    except Exception as e:
        return {"error": str(e)[:200]}


def short_run(torch, capi, name, dev, hbm_peak, fp64_peak, K=5, W=3, dist="ref"):
    """One-line summary of another BASELINE config in the same driver run (device-resident, CUDA events)."""
    w = WORKLOADS[name]
    nx, nu, nc, nth, T, B, legs = w["nx"], w["nu"], w["nc"], w["nth"], w["T"], w["batch"], w["legs"]
    try:
        s = capi.Solver(nx, nu, nc, nth, horizon=T, batch=B, num_legs=legs, device=dev)
        s.generate(seed=42, dist=DIST_ID[dist])
        th = None
        if nth:
            th_t = torch.randn(B, nth, dtype=torch.float64, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
            th = th_t.data_ptr()
        for _ in range(W):
            s.backward(MU, MU); s.forward_device(d_theta=th)
        torch.cuda.synchronize()
        s.timing_log_start(K)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            s.backward(MU, MU); s.forward_device(d_theta=th)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        bwd_ms, _ = s.timing_log_read()
        s.synchronize()
        kkt = s.kkt_error(MU, MU, theta=None if not nth else th_t.cpu().numpy())
        _, f_bwd, _ = flops_model(nx, nu, nc, nx if legs >= 2 else nth, T, nx)
        b_in, b_gain, _ = bytes_model(nx, nu, nc, nx if legs >= 2 else nth, T)
        bwd = float(np.mean(bwd_ms)) * 1e-3
        hf, ff = B * (b_in + b_gain) / bwd / 1e9 / hbm_peak, B * f_bwd / bwd / 1e12 / fp64_peak
        out = {"workload": f"{name}: {w['desc']}", "dist": dist, "explicit_path": bool(s.used_explicit_path()),
               "note": None if dist == "ref" else "fractions use the REFERENCE algorithm's operation count (17 n^3 incl. LU(E), E^-1 products) and record bytes incl. E; the E = -I path executes fewer of both",
               "value": B / (ms * 1e-3), "unit": "solves/s", "ms_per_step": ms,
               "backward_ms": bwd * 1e3, "bound": "hbm" if hf >= ff else "tensor", "roofline_frac": max(hf, ff),
               "hbm_frac": hf, "fp64_frac": ff, "steps": K, "warmup": W,
               "kkt_max": {"dyn": float(np.nanmax(kkt[:, 0])), "cstr": float(np.nanmax(kkt[:, 1])), "dual": float(np.nanmax(kkt[:, 2]))}}
        s.close()
        return out
    except Exception as e:   # a summary row must never take the headline down
        return {"workload": name, "error": str(e)[:200]}


def e2e_run(torch, capi, s_dev, w, K, W, dist, world, theta):
    """Same metric through the C ABI with HOST (pinned) buffers: per step H2D of the packed
    problem, backward, forward, D2H of the trajectories."""
    nx, nu, nc, nth, T, B, legs = w["nx"], w["nu"], w["nc"], w["nth"], w["T"], w["batch"], w["legs"]
    stride = s_dev.download_packed(0, 1)[0].shape[-1]
    per_problem = (T + 1) * stride * 8
    cap = max(1 << 30, (8 << 30) // max(1, world))           # pinned staging: 8 GiB per box, at least 1 GiB per rank
    Be = int(min(B, max(1, cap // per_problem)))
    note = None if Be == B else f"e2e on a {Be}-problem slice per rank (pinned staging capped at {cap / 2**30:.0f} GiB per rank)"

    def pinned(*shape):
        return torch.empty(shape, dtype=torch.float64, pin_memory=True).numpy()

    knots, G0, g0 = pinned(Be, T + 1, stride), pinned(Be, nx * nx), pinned(Be, nx)
    k, a, b = s_dev.download_packed(0, Be)
    knots[:], G0[:], g0[:] = k, a, b
    del k, a, b
    out = (pinned(Be, T + 1, nx), pinned(Be, T + 1, max(nu, 1))[:, :, :nu], pinned(Be, T + 1, max(nc, 1))[:, :, :nc],
           pinned(Be, nx + T * nx))
    if nu == 0 or nc == 0:
        out = tuple(np.ascontiguousarray(o) for o in out)
    th = None if theta is None else np.ascontiguousarray(theta[:Be])

    # The batch goes through NCH independent solver handles (contiguous slices of the problems),
    # each on its own CUDA stream and driven by its own host thread: the H2D copy of one slice
    # overlaps the sweep of another and the D2H of a third (garb200_forward blocks only its own
    # stream; ctypes releases the GIL).  Same C-ABI calls a user would make for a sharded batch.
    from aligator_swk_b200 import sharding
    NCH = 4 if (Be >= 64 and legs < 2) else 1
    dev = torch.cuda.current_device()
    chunks = []
    for c in range(NCH):
        lo, hi = sharding.shard_bounds(Be, c, NCH)
        sc = capi.Solver(nx, nu, nc, nth, horizon=T, batch=hi - lo, num_legs=legs, device=dev)
        st = torch.cuda.Stream()
        sc.set_stream(st.cuda_stream)
        chunks.append((sc, st, lo, hi))

    def run_chunk(c, nsteps):
        sc, st, lo, hi = chunks[c]
        torch.cuda.set_device(dev)
        o = tuple(x[lo:hi] for x in out)
        for _ in range(nsteps):
            sc.upload_packed(knots[lo:hi], G0[lo:hi], g0[lo:hi])
            sc.backward(MU, MU)
            sc.forward(None if th is None else th[lo:hi], out=o)

    def run(nsteps):
        if NCH == 1:
            run_chunk(0, nsteps)
            return
        ts = [threading.Thread(target=run_chunk, args=(c, nsteps)) for c in range(NCH)]
        for t_ in ts:
            t_.start()
        for t_ in ts:
            t_.join()

    run(W)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    torch.cuda.synchronize()
    run(K)                               # every forward() returns with its results in host memory
    torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / K
    h2d = knots.nbytes + G0.nbytes + g0.nbytes + (th.nbytes if th is not None else 0)
    d2h = sum(o.nbytes for o in out)
    r = {"value": Be * world / (ms * 1e-3), "unit": "solves/s", "h2d_bytes_per_step": int(h2d),
         "d2h_bytes_per_step": int(d2h), "ms_per_step": ms, "batch_per_gpu": Be}
    r["streams"] = NCH
    if note:
        r["note"] = note
    for sc, _, _, _ in chunks:
        sc.close()
    return r


def horizon_sweep(args):
    """The reference's own benchmark (bench/gar-riccati.cpp:18-31,62-68): ONE random problem nx=36 nu=12,
    horizons 16 ... 1024 (+ T = 100, tests/gar/riccati.cpp:113), backward + forward latency.  GPU: the serial
    sweep and the parallel-in-time solver (legs ~ 0.83 sqrt(T), the latency optimum of leg time vs condensed solve), CUDA events, mean over repetitions;
    CPU: the oracle port, serial (1 thread) and parallel legs (all host cores), next to it."""
    import torch
    import garb200_pkg
    garb200_pkg.load()
    from aligator_swk_b200 import capi
    from aligator_swk_b200.synth import generate_problems
    import pyoracle as po
    po.build()
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    nx, nu = 36, 12
    rows = []
    # leg time ~ 2 x 13 us x T / J, condensed solve ~ 2 J x 19 us  ->  J* ~ 0.83 sqrt(T)
    gpu_legs = lambda T: int(min(64, max(2, round(0.83 * T ** 0.5))))
    for T in (16, 32, 64, 100, 128, 256, 512, 1024):
        pp = generate_problems(1, T, nx, nu, 0, 0, seed=42)
        row = {"horizon": T}
        for name, legs in (("gpu_serial_ms", 0), ("gpu_legs_ms", gpu_legs(T))):
            s = capi.Solver.for_problems(pp, num_legs=legs)
            s.upload(pp)
            for _ in range(3):
                s.backward(MU, MU); s.forward_device()
            torch.cuda.synchronize()
            reps = 20
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                s.backward(MU, MU); s.forward_device()
            e1.record(); torch.cuda.synchronize()
            s.synchronize()
            row[name] = e0.elapsed_time(e1) / reps
            if legs:
                row["legs"] = legs
            row[name.replace("_ms", "_kkt")] = float(s.kkt_error(MU, MU).max())
            s.close()
        for name, legs, nthreads in (("cpu_serial_ms", 0, 1), ("cpu_legs_ms", min(cores, max(2, T // 4)), cores)):
            secs, _ = po.bench(pp, MU, MU, num_legs=legs, reps=1, nthreads=nthreads)
            reps = int(max(3, min(50, 2.0 / max(secs, 1e-6))))
            secs, used = po.bench(pp, MU, MU, num_legs=legs, reps=reps, nthreads=nthreads)
            row[name] = secs / reps * 1e3
            if legs:
                row["cpu_legs"], row["cpu_threads"] = legs, used
        rows.append(row)
        print("T=%4d  gpu serial %.3f ms  gpu %d legs %.3f ms | cpu serial %.2f ms  cpu %d legs %.2f ms" % (
            T, row["gpu_serial_ms"], row["legs"], row["gpu_legs_ms"], row["cpu_serial_ms"], row["cpu_legs"], row["cpu_legs_ms"]),
            file=sys.stderr, flush=True)
    print(json.dumps({"metric": "single-problem backward+forward latency over horizons (bench/gar-riccati.cpp)", "unit": "ms",
                      "higher_is_better": False, "nx": nx, "nu": nu, "mu": MU, "cpu_kind": "port", "cpu_cores": cores,
                      "rows": rows}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c1b", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="override the per-GPU batch")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU-port timing")
    ap.add_argument("--parity-k", type=int, default=8, help="problems per rank re-solved by the CPU oracle after the timed region (0: off)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank gets the config's batch; strong: the config's batch is split over the ranks")
    ap.add_argument("--gather", default="nccl", choices=["auto", "p2p", "nccl"],
                    help="trajectory gather to rank 0: NCCL (default, the north_star's single gather) or peer copies through a "
                         "CUDA IPC mapping (no SMs; measured on 2 B200s: no gain on c1b, 6.4 vs 3.8 ms/step on c2 — torch's "
                         "cross-process peer copy does not overlap like the NCCL kernel does); auto tries p2p first")
    ap.add_argument("--dist", default="ref", choices=["ref", "explicit", "explicit_psd"], help="synthetic distribution (SURVEY 8d)")
    ap.add_argument("--others", default="c2,c3,c5,c1b+explicit_psd,c2+explicit_psd,c5+explicit_psd,c1b+explicit,p30", help="other BASELINE configs summarised in other_configs at N = 1 ('' = none)")
    ap.add_argument("--tune", action="append", default=[], help="diagnostic launcher knob key=value (garb200_set_tuning)")
    ap.add_argument("--horizons", action="store_true", help="the reference's own latency sweep over horizons (one GPU)")
    args = ap.parse_args()
    args.tune_knobs = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in args.tune}
    if args.horizons:
        horizon_sweep(args)
        return
    w = dict(WORKLOADS[args.workload])
    if args.batch > 0:
        w["batch"] = args.batch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        # not launched by torchrun: re-launch ourselves one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 2000), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        reference_arm(args, w, rank, world)
        return
    gpu_arm(args, w, rank, world, local_rank)


if __name__ == "__main__":
    main()
