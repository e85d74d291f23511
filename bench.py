#!/usr/bin/env python
"""bench.py — BA observations/s per LM (trust-region) iteration on B200, with the RANSAC hypotheses/s leg,
the residual/Jacobian kernel's HBM roofline and the reference's CPU path timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg4shard|cfg1|cfg3] [--impl reference]

One "step" = one pass of SciPy's trust-region outer loop body on the whole scene: damped Schur solves until a
step is accepted (each with its cost evaluation) and one re-linearisation.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[3]: 64 cameras x 100k frames x 17 joints sharded by frame over 8 GPUs; one GPU's
    # shard is 12 500 frames, so N GPUs hold N x 12 500 frames (weak scaling; N = 8 is the full config).
    "cfg4shard": dict(C=64, N=12500, kind="cfg4", lam=(1.0, 1.0),
                      name="cfg4 shard: 64 cameras x 12500 frames/GPU x 17 joints (8 GPUs = BASELINE configs[3], 1.088e8 obs)"),
    # the whole of configs[3] on ONE GPU (it fits: observations 2.6 GB, whitened blocks 16 GB); not the default because
    # the N-GPU runs are weak-scaling over shards and 8 shards make exactly this scene
    "cfg4full": dict(C=64, N=100000, kind="cfg4", lam=(1.0, 1.0),
                     name="cfg4 full: 64 cameras x 100000 frames x 17 joints on one GPU (BASELINE configs[3], 1.088e8 obs)"),
    "cfg3": dict(C=31, N=10000, kind="cfg3", lam=(100.0, 1.0),
                 name="cfg3: Panoptic-shaped, 31 cameras x 10000 frames x 17 joints, 30% dropped detections (per GPU)"),
    "cfg1": dict(C=4, N=1000, kind="cfg1", lam=(1.0, 1.0),
                 name="cfg1: SynADL-shaped, 4 cameras x 1000 frames x 17 joints (per GPU)"),
}


def make_workload(name, seed):
    import calib_b200  # noqa: F401
    from calib_b200 import scene
    w = WORKLOADS[name]
    if w["kind"] == "cfg4":
        sc = scene.cfg4(seed=seed, N=w["N"], C=w["C"])
    elif w["kind"] == "cfg3":
        sc = scene.cfg3(seed=seed, N=w["N"], C=w["C"])
    else:
        sc = scene.cfg1(seed=seed, N=w["N"])
    # frames differ per rank (seed), the camera start is common to all ranks (replicated parameters)
    common = scene.make_scene(C=w["C"], N=2, seed=999)
    sc.R_init, sc.t_init = common.R_init, common.t_init
    return sc, scene.ba_inputs(sc), w


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 7:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def bind_to_gpu_numa_node(local_rank):
    """Best effort: run this rank on the CPUs of the NUMA node its GPU hangs off BEFORE the pinned observation buffers
    are allocated, so that first touch puts them in node-local DRAM (8 ranks x 331 MB of upload per step otherwise cross
    the socket interconnect: VERDICT r01, e2e at 4 / 8 GPUs).  Returns a short description for the bench line."""
    try:
        import torch
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = vis.split(",")[local_rank].strip() if vis else str(local_rank)      # nvidia-smi counts physical devices
        out = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", phys],
                             capture_output=True, text=True, timeout=10).stdout.strip()
        bus = out.splitlines()[0].strip().lower()
        if bus.count(":") == 2 and len(bus.split(":")[0]) == 8:
            bus = bus[4:]                                  # 00000000:1B:00.0 -> 0000:1b:00.0
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read().strip())
        if node < 0:
            return "numa: single node / unknown"
        cpus = []
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.extend(range(int(a), int(b or a) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return f"numa: node {node} has no allowed CPU"
        os.sched_setaffinity(0, allowed)
        return f"numa: rank bound to node {node} ({len(allowed)} CPUs)"
    except Exception as e:          # no sysfs / no permission: leave the affinity alone
        return f"numa: not bound ({type(e).__name__})"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


_CPU_JOB = {}
REF_PROCS = 8


def _cpu_worker(reps):
    """One host process of the reference arm: `reps` calls of the oracle port of ba.objfun on the forked inputs."""
    from oracle import ba_oracle as bo
    out = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f = bo.objfun(_CPU_JOB["theta"], *_CPU_JOB["args"])
        out.append(time.perf_counter() - t0)
    return out, int(f.size)


def cpu_reference_leg(sc, inputs, w, budget_s=20.0, max_calls=8, procs=None):
    """The reference's CPU algorithm for one LM iteration (ba.py:187-206 via SciPy TRF): a dense 2-point
    finite-difference Jacobian = n + 1 calls of ba.objfun, then a dense SVD.  The n + 1 calls are independent, so the
    arm gives them every host core: `procs` forked processes each time calls of the oracle port of ba.objfun at the
    full workload size (a bounded sample), and the per-iteration cost is the extrapolation (n + 1) / (calls per
    second over all processes); the SVD and the trial evaluations are left out, in the reference's favour.  (The
    reference itself is single-threaded Python: SciPy's approx_derivative walks the columns in one thread.)
    Must run in a process that has not initialised CUDA (fork)."""
    import multiprocessing as mp
    from oracle import ba_oracle as bo
    sp2d, ss2d, sp3d, ss3d = inputs
    C, N, J = ss2d.shape
    theta = np.concatenate([np.zeros(3 * C), sc.t_init.reshape(-1), sc.X_w[:, :J].reshape(-1)])
    theta[:3 * C] = np.concatenate([bo.rodrigues_inv(R) for R in sc.R_init])
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, w["lam"][0], w["lam"][1])
    _CPU_JOB["theta"], _CPU_JOB["args"] = theta, args
    t0 = time.perf_counter()
    f = bo.objfun(theta, *args)                                   # one call alone: single-core figure + sizing
    t_one = time.perf_counter() - t0
    # a FIXED process count (not os.cpu_count()): the r01 arm forked one worker per core and its calls/s swung 9x
    # between boxes / co-tenants as 32 workers fought for host DRAM
    procs = int(procs or min(REF_PROCS, os.cpu_count() or 1))
    try:                                                          # stay inside the host memory: ~40 B x m per call
        import psutil
        procs = max(1, min(procs, int(psutil.virtual_memory().available * 0.5 / max(1.0, 60.0 * f.size * 8))))
    except Exception:
        pass
    reps = int(max(1, min(max_calls, budget_s / max(t_one * 1.5, 1e-3))))
    t0 = time.perf_counter()
    if procs > 1:
        with mp.get_context("fork").Pool(procs) as pool:
            res = pool.map(_cpu_worker, [reps] * procs)
    else:
        res = [_cpu_worker(reps)]
    wall = time.perf_counter() - t0
    times = [t for r in res for t in r[0]]
    calls_per_s = procs * reps / wall
    n_obs = int((ss2d > 0).sum())
    n_par = theta.size
    per_iter = (n_par + 1) / calls_per_s
    return dict(value=n_obs / per_iter, unit="observations/s per LM iteration", cores=procs, kind="port",
                sample=f"{procs} processes x {reps} calls of the oracle port of ba.objfun at full workload size in {wall:.1f} s "
                       f"({calls_per_s:.2f} calls/s; one call alone {t_one * 1e3:.0f} ms, m={f.size}); one reference iteration = "
                       f"n+1 = {n_par + 1} such calls (dense 2-point FD Jacobian, ba.py:187-206) -> extrapolated "
                       f"{per_iter:.3g} s/iteration on {procs} cores ({(n_par + 1) * t_one:.3g} s on 1); dense SVD of the "
                       f"{f.size}x{n_par} Jacobian not included",
                t_objfun_ms=t_one * 1e3, calls_per_s=calls_per_s, single_core_value=n_obs / ((n_par + 1) * t_one),
                call_s_min_median_max=[float(np.min(times)), float(np.median(times)), float(np.max(times))],
                n_obs=n_obs, host_cpus=os.cpu_count()), times


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the Python reference cannot travel to the GPU
    box) on this arm's config / metric / unit.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sc, inputs, w = make_workload(args.workload, 1000)
    t0 = time.perf_counter()
    leg, times = cpu_reference_leg(sc, inputs, w, budget_s=30.0, max_calls=max(1, args.steps))
    if not args.no_cpu_rows:
        # EXECUTED CPU rows (not extrapolated): the literal reference algorithm at frame counts it can run, and the
        # strong non-reference row (analytic sparse Jacobian + LSMR) at cfg1 size
        leg["executed"] = cpu_rows(budget_s=25.0)
    line = {"impl": "reference", "metric": "BA observations/s per LM iteration", "value": leg["value"],
            "unit": "observations/s", "n_gpus": args.gpus, "steps": len(times) // leg["cores"], "warmup": 1,
            "ms_per_step": float(np.mean(times)) * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["name"], "cameras": w["C"], "frames_per_gpu": w["N"], "joints": 17,
                       "step": "bounded sample: one ba.objfun call per host process; value extrapolates the n+1 calls of one FD "
                               "Jacobian spread over all host cores"},
            "cpu_baseline": leg,
            "e2e": {"value": leg["value"], "unit": "observations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - t0}
    print(json.dumps(line))


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import calib_b200  # noqa: F401
    from calib_b200 import ba

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD

    sc, (sp2d, ss2d, sp3d, ss3d), w = make_workload(args.workload, 1000 + rank)
    C, N, J = ss2d.shape
    lam1, lam2 = w["lam"]
    cpus0 = os.sched_getaffinity(0)
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else "numa: one GPU, not bound"
    y_h = torch.from_numpy(sp2d.reshape(C, N * J, 2)).pin_memory()     # first touch on the GPU's NUMA node
    s_h = torch.from_numpy(ss2d.reshape(C, N * J)).pin_memory()
    os.sched_setaffinity(0, cpus0)                                     # the worker threads of the later legs roam again
    prob = ba.BAProblem(sc.K, y_h, s_h.reshape(C, N, J), sp3d, ss3d, ba.OP_BONE, lam1, lam2, device=dev,
                        group=group, n_frames_global=N * world)
    x0 = prob.triangulate(sc.R_init, sc.t_init)
    theta0 = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    theta_h = torch.from_numpy(theta0.copy()).pin_memory()
    n_obs_local = prob.n_vis
    n_obs = n_obs_local
    if world > 1:
        t = torch.tensor([n_obs_local], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        n_obs = int(t.item())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world > 1:
            t = torch.tensor([v], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return v

    # ---------------- device-resident LM iterations
    class Runner:
        """LM iterations of complete solves: iterate from theta0 with the reference's own tolerances (ftol 1e-4,
        xtol 1e-8, gtol 1e-8: ba.py:187-206) and, when SciPy's termination test fires, start the next solve from
        theta0 — so the timed steps are the iteration mix of real runs (first iterations included), not iterations
        at a stationary point.  The restart's own linearisation is inside the timed region."""

        def __init__(self):
            self.nfev, self.restarts = 0, 0
            self.sol = ba.TRFSolver(prob, theta0)

        def iterate(self):
            self.sol.iterate()
            if self.sol.done():
                self.nfev += self.sol.nfev
                self.restarts += 1
                self.sol = ba.TRFSolver(prob, theta0)

        def counts(self):
            return (self.nfev + self.sol.nfev, prob.n_trials)

    run = Runner()
    for _ in range(args.warmup):
        run.iterate()
    launches0 = sum(v[1] for v in prob.timing_read().values())
    nfev0, trials0 = run.counts()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        run.iterate()
    ev1.record()
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    sampler.stop_flag = True
    launches = sum(v[1] for v in prob.timing_read().values()) - launches0
    nfev1, trials1 = run.counts()
    nfev, trials = nfev1 - nfev0, trials1 - trials0
    sol = run.sol
    value = n_obs * args.steps / (ms_total * 1e-3)

    # ---------------- end to end: host buffers in, host result out, every step
    lm_cost = sol.cost
    theta0_h = torch.from_numpy(theta0.copy()).pin_memory()
    st8 = {"Delta": None, "alpha": 0.0}
    theta_h.copy_(theta0_h)
    torch.cuda.synchronize()
    x_dev = torch.empty_like(sol.x)

    # double-buffered observation upload: while step k computes on buffer k % 2, the copy stream uploads the pinned
    # host observations of step k+1 into the other buffer (every step still copies its own inputs host -> device)
    ybuf = [prob.y, torch.empty_like(prob.y)]
    sbuf = [prob.s, torch.empty_like(prob.s)]
    copy_stream = torch.cuda.Stream(device=dev)
    up_done = [torch.cuda.Event(), torch.cuda.Event()]
    free_ev = [torch.cuda.Event(), torch.cuda.Event()]
    st8["k"] = 0

    def upload(i):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(free_ev[i])         # kernels of the step that last read buffer i are done
            ybuf[i].copy_(y_h, non_blocking=True)
            sbuf[i].copy_(s_h, non_blocking=True)
            up_done[i].record(copy_stream)

    for i in range(2):
        free_ev[i].record()
    upload(0)

    def e2e_step():
        i = st8["k"] % 2
        st8["k"] += 1
        x_dev.copy_(theta_h, non_blocking=True)        # first: the copy engine serves requests in issue order
        upload(1 - i)                                  # next step's inputs, overlapped with this step's kernels
        torch.cuda.current_stream().wait_event(up_done[i])
        prob.rebind_observations(ybuf[i], sbuf[i])
        s2 = ba.TRFSolver(prob, x_dev)                 # linearises at the uploaded theta
        if st8["Delta"] is not None:
            s2.Delta, s2.alpha = st8["Delta"], st8["alpha"]
        s2.iterate(relinearize=False)
        free_ev[i].record()
        theta_h.copy_(s2.x, non_blocking=True)
        torch.cuda.current_stream().synchronize()      # result on the host; the next upload keeps running
        if s2.status is None:
            st8["Delta"], st8["alpha"] = s2.Delta, s2.alpha
        else:                                          # solve finished: the next step starts a new one
            theta_h.copy_(theta0_h)
            st8["Delta"], st8["alpha"] = None, 0.0
        return s2

    for _ in range(min(args.warmup, 2)):
        e2e_step()
    k_e2e = max(1, args.steps)
    barrier()
    t0 = time.perf_counter()
    for _ in range(k_e2e):
        e2e_step()
    torch.cuda.synchronize()
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    e2e_value = n_obs * k_e2e / t_e2e
    h2d = y_h.numel() * 8 + s_h.numel() * 8 + theta_h.numel() * 8
    d2h = theta_h.numel() * 8

    # ---------------- residual/Jacobian kernel against the HBM roofline (measured live, CUDA events)
    peak, peak_src = measured_peaks()
    out = prob.residual_jacobian(sol.x, with_index=False)
    for _ in range(3):
        prob.residual_jacobian(sol.x, with_index=False, out=out)
    n_rj = 10
    barrier()
    ev0.record()
    for _ in range(n_rj):
        prob.residual_jacobian(sol.x, with_index=False, out=out)
    ev1.record()
    torch.cuda.synchronize()
    ms_rj = ev0.elapsed_time(ev1) / n_rj          # includes the 1-block camera-pack kernel (~2 us)
    cbar = n_obs_local / float(prob.P)
    bytes_alg = (184.0 + 24.0 / cbar) * n_obs_local
    achieved = bytes_alg / (ms_rj * 1e-3) / 1e9
    del out
    roofline = {"kernel": "k_rj (fused residual+Jacobian, bax_residual_jacobian)", "bound": "hbm",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_obs": 184.0 + 24.0 / cbar,
                "in_timed_step": False,     # the LM step linearises with fused-accumulate kernels and never materialises J
                "ms_per_launch": ms_rj, "obs_per_launch": n_obs_local, "obs_per_s": n_obs_local / (ms_rj * 1e-3)}
    prof = os.path.join(ROOT, "profiles", "rj_traffic.json")
    if os.path.isfile(prof):
        with open(prof) as f:
            tr = json.load(f)
        if tr.get("obs_per_launch") == n_obs_local:
            roofline["traffic"] = tr.get("dram_bytes_per_launch")

    # ---------------- kernel shares of one LM iteration (separate instrumented pass; events add syncs)
    sol = ba.TRFSolver(prob, theta0, ftol=0.0, xtol=0.0, gtol=0.0)
    prob.timing(True)
    for _ in range(2):
        sol.iterate()
    torch.cuda.synchronize()
    shares = {k: {"ms": round(v[0], 4), "launches": v[1]} for k, v in prob.timing_read().items() if v[1]}
    prob.timing(False)
    # ---------------- the kernel that bounds the timed step: k_syrk96 (FP64 DMMA SYRK of the whitened Schur blocks)
    roofline_step = syrk_roofline(prob, sol, ms_total / args.steps, C)
    parity = sharded_parity_probe(rank, world, dev, group)
    comm = None
    if world > 1:      # device time inside the all-reduces (waiting for the slowest rank included), separate pass
        names = ["bone", "cost", "lin", "schur", "rhs2", "step", "norm", "gmax"]
        prob.comm_events = []       # instrumented pass: the host-language loop with timed torch.distributed all-reduces
        sol = ba.TRFSolver(prob, theta0, ftol=0.0, xtol=0.0, gtol=0.0)
        barrier()
        ev0.record()
        for _ in range(6):
            sol.iterate()
        ev1.record()
        torch.cuda.synchronize()
        comm = {names[k]: {"calls": v[0], "ms": round(v[1], 4), "doubles": v[2]} for k, v in prob.comm_timing_read().items()}
        comm["ms_6_steps"] = ev0.elapsed_time(ev1)
        prob.comm_events = None

    line = None
    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            # a fresh process (this one holds a CUDA context; the leg forks one worker per host core)
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--workload", args.workload,
                                  "--steps", "3", "--warmup", "1"], capture_output=True, text=True)
            try:
                cpu = json.loads(out.stdout.strip().splitlines()[-1])["cpu_baseline"]
            except Exception:
                raise RuntimeError("cpu_baseline leg failed: " + out.stderr[-2000:])
        line = {"metric": "BA observations/s per LM iteration", "value": value, "unit": "observations/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": w["name"], "cameras": C, "frames_per_gpu": N, "joints": J,
                           "observations_total": n_obs, "parameters_per_gpu": prob.n,
                           "lambda1": lam1, "lambda2": lam2,
                           "step": "one trust-region outer iteration (SciPy trf_no_bounds body): damped Schur solves + "
                                   "cost evaluations until acceptance, then re-linearisation",
                           "cache": "inputs larger than L2 (observations + whitened blocks >> 126 MB)",
                           "parallelism": f"frames sharded over {world} GPU(s); all-reduce of the packed camera blocks",
                           "collectives": ("NCCL issued inside libbax between the kernel phases (5 per iteration)"
                                           if getattr(prob, "native_comm", None) is not None else
                                           ("torch.distributed between ctypes calls" if world > 1 else "none (one GPU)"))},
                "lm": {"cost_evaluations_per_step": nfev / args.steps, "damped_solves_per_step": trials / args.steps,
                       "cost": lm_cost, "restarts": run.restarts, "kernel_ms_2_steps": shares},
                "e2e": {"value": e2e_value, "unit": "observations/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "steps": k_e2e, "ms_per_step": t_e2e / k_e2e * 1e3,
                        "h2d_gb_per_s_per_gpu": h2d / (t_e2e / k_e2e) / 1e9, "host_placement": numa,
                        "note": "every step: pinned-host observations + theta -> device, one full LM iteration "
                                "(linearise, solves, cost evaluations), theta -> host; same solve-restart rule as value; the upload of step "
                                "k+1 (second device buffer, copy stream) overlaps the kernels of step k"},
                "gpu_launches": int(launches), "roofline": roofline, "roofline_step": roofline_step,
                "sharded_parity": parity, "cpu_baseline": cpu, "comm": comm,
                "clocks": sampler.summary()}
    # ---------------- the other single-GPU BA configs, short runs (BASELINE configs[0] and [2]); one GPU only
    if line is not None and world == 1 and not args.no_other:
        line["other_configs"] = {name: other_config_leg(name, dev) for name in ("cfg1", "cfg3") if name != args.workload}
    # ---------------- RANSAC leg (hypotheses/s), BASELINE configs[1]
    from calib_b200 import calib_ransac as cr
    rs = cr.bench_leg(rank, world, dev, steps=max(1, min(args.steps, 3)), no_cpu=args.no_cpu or world > 1)
    cpu_in = rs.pop("_cpu_inputs", None)
    if cpu_in is not None:
        rs["cpu_baseline"] = ransac_cpu_leg(**cpu_in)
    if line is not None:
        line["ransac"] = rs
        line["gpu_launches"] = int(launches) + sum(v["launches"] for v in rs["kernel_ms_one_pass"].values()) * max(1, min(args.steps, 3))
    # ---------------- sweep leg, BASELINE configs[4]: independent cfg1-shaped problems (a bounded sample per GPU)
    bt = batch_leg(rank, world, group, n_per_gpu=args.batch)
    if line is not None and bt is not None:
        line["batch"] = bt
    if rank == 0:
        print(json.dumps(line))
    prob.close()
    if world > 1:
        dist.destroy_process_group()


def other_config_leg(name, dev, steps=30, warmup=3):
    """BA observations/s per LM iteration on another BASELINE config (device-resident inputs, CUDA events, iterations of
    complete solves with the reference's tolerances, as for the headline workload) + its k_rj fraction of the HBM peak."""
    import torch
    from calib_b200 import ba
    sc, (sp2d, ss2d, sp3d, ss3d), w = make_workload(name, 1000)
    C, N, J = ss2d.shape
    prob = ba.BAProblem(sc.K, sp2d, ss2d, sp3d, ss3d, ba.OP_BONE, *w["lam"], device=dev)
    try:
        x0 = prob.triangulate(sc.R_init, sc.t_init)
        theta0 = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
        sol = ba.TRFSolver(prob, theta0)
        its = 0

        def step():
            nonlocal sol
            sol.iterate()
            if sol.done():
                sol = ba.TRFSolver(prob, theta0)
        for _ in range(warmup):
            step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out = prob.residual_jacobian(sol.x, with_index=False)
        for _ in range(3):
            prob.residual_jacobian(sol.x, with_index=False, out=out)
        e0.record()
        for _ in range(10):
            prob.residual_jacobian(sol.x, with_index=False, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms_rj = e0.elapsed_time(e1) / 10
        peak, _ = measured_peaks()
        bpo = 184.0 + 24.0 / (prob.n_vis / float(prob.P))
        return {"workload": w["name"], "observations": prob.n_vis, "value": prob.n_vis / (ms * 1e-3), "unit": "observations/s",
                "ms_per_step": ms, "steps": steps, "k_rj_frac_of_hbm_peak": bpo * prob.n_vis / (ms_rj * 1e-3) / 1e9 / peak,
                "k_rj_ms": ms_rj}
    finally:
        prob.close()


def fp64_peaks():
    """FP64 peaks measured on this pool's B200 by tools/micro/fp64_peak.cu (tools/fp64_peak_json.py)."""
    for rnd in ("r02",):
        p = os.path.join(ROOT, "profiles", rnd, "fp64_peak.json")
        if os.path.isfile(p):
            with open(p) as f:
                d = json.load(f)
            return float(d["dmma_tflops"]), float(d["dfma_tflops"]), f"measured (profiles/{rnd}/fp64_peak.json)"
    return 37.0, 34.0, "fallback (B200 datasheet FP64)"


def syrk_roofline(prob, sol, ms_step, C):
    """k_syrk96 against the FP64 tensor (DMMA) peak: timed live with CUDA events (the library's own per-family event
    timers) over damped solves at the current linearisation, so the bucket holds only the trial SYRK (k_syrk96 + its two
    split-K reduction kernels, microseconds).  Algorithmic flops = the lower triangle incl. diagonal of the
    (6C+B+1)^2 product over K = 3P columns; issued = what the 96x96 tiles execute: 144 m8n8 accumulator tiles per
    off-diagonal tile, and per diagonal tile the 78 of its lower triangle + 24 of the extras panel (+ 4 for the extras
    corner in tile 0) — padding rows and the upper halves of the 8x8 tiles on the diagonal included."""
    import torch
    dmma, dfma, src = fp64_peaks()
    n_red = 6 * C + 12 + 1
    K = 3 * prob.P
    nt = (6 * C + 95) // 96
    tiles = nt * (nt + 1) // 2
    mma_tiles = (tiles - nt) * 144 + nt * (78 + 24) + 4
    issued = mma_tiles * 64.0 * K * 2.0
    algorithmic = n_red * (n_red + 1) / 2.0 * K * 2.0
    alpha = 0.01 * sol.g_2 / max(sol.x_norm, 1e-300)
    prob.timing(True)
    reps = 6
    for _ in range(reps):
        prob.trial(alpha, False)
    torch.cuda.synchronize()
    t = prob.timing_read()
    prob.timing(False)
    ms = t["syrk"][0] / reps
    return {"kernel": "k_syrk96 (FP64 DMMA SYRK of the whitened Schur blocks, one per damped solve)", "bound": "tensor",
            "pipe": "fp64 mma.sync.m8n8k4 (tcgen05 has no FP64 kind)", "achieved": algorithmic / (ms * 1e-3) / 1e12,
            "peak": dmma, "unit": "TFLOP/s", "frac": algorithmic / (ms * 1e-3) / 1e12 / dmma, "traffic": None,
            "issued_tflops": issued / (ms * 1e-3) / 1e12, "issued_frac": issued / (ms * 1e-3) / 1e12 / dmma,
            "dfma_peak": dfma, "peak_source": src, "flops_per_launch_algorithmic": algorithmic,
            "flops_per_launch_issued": issued, "tiles_96x96": tiles, "diagonal_tiles": nt, "mma_tiles_m8n8": mma_tiles, "k_columns": K, "reduced_size": n_red,
            "ms_per_launch": ms, "launches_per_solve": t["syrk"][1] // reps, "share_of_step": ms / ms_step,
            "in_timed_step": True}


def sharded_parity_probe(rank, world, dev, group, C=16, frames=960, seed=4242):
    """The SAME fixed scene and start on every N: its frames are sharded over the N ranks, the cost and gradient norm at
    theta0 and the cost after two LM iterations are reported — they must agree between the N = 1, 2, 4, 8 lines of the
    driver's scaling run (sharded-parity check riding on SCALE)."""
    import calib_b200  # noqa: F401
    from calib_b200 import ba, scene
    sc = scene.make_scene(C=C, N=frames, seed=seed, score_mode="uniform", p_drop_2d=0.2, p_drop_3d=0.1)
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    per = frames // world
    a, b = rank * per, (rank + 1) * per if rank < world - 1 else frames
    prob = ba.BAProblem(sc.K, np.ascontiguousarray(sp2d[:, a:b]), np.ascontiguousarray(ss2d[:, a:b]),
                        np.ascontiguousarray(sp3d[:, a:b]), np.ascontiguousarray(ss3d[:, a:b]), ba.OP_BONE, 3.0, 7.0,
                        device=dev, group=group, n_frames_global=frames)
    try:
        x0 = prob.triangulate(sc.R_init, sc.t_init)
        theta0 = ba.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
        sol = ba.TRFSolver(prob, theta0, ftol=0.0, xtol=0.0, gtol=0.0)
        cost0, g0 = sol.cost, sol.g_2
        alpha = 0.01 * sol.g_2 / sol.x_norm
        pn2, gp, ptq = prob.trial(alpha, True)                       # one damped solve at a fixed, sizeable alpha
        # two iterations inside a SMALL trust region (boundary steps, alpha well above its floor): a step at the alpha
        # floor is dominated, in the gauge null space of the exact Jacobian, by round-off / alpha and differs between
        # summation orders (shard counts) while staying a valid step — the cost after such steps agrees to ~1e-5 only
        sol.Delta = 1e-3 * sol.x_norm
        sol.iterate(); sol.iterate()
        out = {"scene": f"{C} cameras x {frames} frames x 17 joints, seed {seed}, lambda 3/7, frames sharded over {world} GPU(s)",
               "cost0": cost0, "g_norm0": g0, "damped_step_alpha": alpha, "damped_step_norm2": pn2, "damped_step_g_dot_p": gp,
               "damped_step_phi_prime_term": ptq, "cost_after_2_boundary_iterations": sol.cost, "nfev": sol.nfev,
               "must_agree_over_n": "every number here, to ~1e-12 relative (different summation orders only)"}
    finally:
        prob.close()
    return out


def ransac_cpu_leg(v, n, sn, sm, scores, budget_s=15.0):
    """cpu_baseline of the RANSAC leg: the first hypotheses of the same sample stream through the oracle port on one
    host core (bounded to ~15 s), their scores compared with the GPU's."""
    from oracle import ransac_oracle as ro
    k, t2, worst = 0, time.perf_counter(), 0.0
    while k < len(sn) and (time.perf_counter() - t2 < budget_s or k < 2):
        Rh, th_, _ = ro.calib_linear(v[:, sn[k]], n[:, sm[k]])
        _, _, es, e3s = ro.score_hypothesis(Rh, th_, v, n)
        worst = max(worst, abs(es + e3s - scores[k]) / (es + e3s))
        k += 1
    dt = time.perf_counter() - t2
    return {"value": k / dt, "unit": "hypotheses/s", "cores": 1, "kind": "port",
            "sample": f"first {k} hypotheses of the same seed at full M, N_o through the vectorised oracle port "
                      f"(the literal reference loop is ~20x slower, SURVEY.md §6)",
            "max_rel_score_diff_vs_gpu": worst}


def batch_leg(rank, world, group, n_per_gpu=512, n_streams=12, n_distinct=64):
    """BASELINE configs[4]: 4096 independent 4-camera x 1000-frame problems = 512 per GPU on 8 GPUs, so every GPU
    solves n_per_gpu = 512 problems (weak scaling; the 8-GPU line is the full config).  To keep the bench's host memory
    and scene-generation time bounded, a GPU's problems are n_distinct = 64 different seeded scenes, each solved
    n_per_gpu / n_distinct times (the work of a problem does not depend on whether its data was seen before: every
    problem is uploaded, bound, triangulated and solved from scratch).  Every problem is a complete ba_main run
    (triangulated start, reference tolerances), host arrays in and host results out, through ba.ba_batch ->
    bax_batch_solve (native worker threads, one stream + handle each)."""
    if n_per_gpu <= 0:
        return None
    import torch
    import torch.distributed as dist
    from calib_b200 import ba, scene
    n_distinct = max(1, min(n_distinct, n_per_gpu))
    n = n_per_gpu * world
    scenes = {}
    probs = []
    for k in range(n):
        if k % world != rank:
            probs.append(None)            # never touched on this rank
            continue
        d = (k // world) % n_distinct
        if d not in scenes:
            sc = scene.cfg1(seed=7000 + d * world + rank, N=1000)
            scenes[d] = (sc.camid, sc.K, sc.R_init, sc.t_init) + tuple(scene.ba_inputs(sc))
        probs.append(scenes[d])
    mine = [probs[k] for k in ba.batch_plan(n, rank, world)]
    ba.ba_batch(mine[:min(len(mine), n_streams)], 1.0, 1.0, n_streams=n_streams)         # warm-up

    def run(ps, ns, grp):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = ba.ba_batch(ps, 1.0, 1.0, n_streams=ns, group=grp, gather=False, return_info=True)
        torch.cuda.synchronize()
        return time.perf_counter() - t0, out
    t_multi, out = run(probs, n_streams, group)
    work = float(sum(int((np.asarray(p[5]) > 0).sum()) * r[3]["iterations"] for p, r in zip(mine, out)))
    iters = float(sum(r[3]["iterations"] for r in out))
    if world > 1:
        v = torch.tensor([t_multi], dtype=torch.float64, device="cuda")
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        t_multi = float(v[0])
        s = torch.tensor([work, iters], dtype=torch.float64, device="cuda")
        dist.all_reduce(s)
        work, iters = float(s[0]), float(s[1])
    res = {"metric": "BA problems/s (complete ba_main runs, host arrays in / out)", "value": n / t_multi, "unit": "problems/s",
           "problems": n, "problems_per_gpu": n_per_gpu, "distinct_scenes_per_gpu": n_distinct,
           "workers_per_gpu": ba.batch_stats.get("workers", n_streams), "workers_requested": n_streams,
           "host": {k: ba.batch_stats.get(k) for k in ("cpus", "local_world", "blocking_sync")},
           "wall_s": t_multi, "lm_iterations": iters, "observations_per_s_per_lm_iteration": work / t_multi,
           "config": {"workload": "cfg5: independent 4-camera x 1000-frame x 17-joint problems, 512 per GPU (8 GPUs = the 4096 "
                                  "of BASELINE configs[4])",
                      "sharding": f"problems dealt round-robin over {world} GPU(s); replicas only, no data-path collective"}}
    if rank == 0:
        t_single, _ = run_local_single(mine[:n_distinct], ba)
        res["single_worker_problems_per_s_one_gpu"] = min(len(mine), n_distinct) / t_single
    return res


def run_local_single(mine, ba):
    import torch
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = ba.ba_batch(mine, 1.0, 1.0, n_streams=1)
    torch.cuda.synchronize()
    return time.perf_counter() - t0, out


def run_cpu_rows(args):
    print(json.dumps(cpu_rows(float(os.environ.get("CPU_ROWS_BUDGET_S", "60")))))


def cpu_rows(budget_s=60.0):
    """`--impl cpu_rows` and the `executed` object of cpu_baseline: the CPU rows that stand beside the GPU numbers, on this box's host
    cores.  literal: SciPy least_squares(trf, dense 2-point finite differences, dense SVD) on the oracle port of ba.objfun —
    what reference ba.py:187-206 runs — at frame counts small enough to execute.  strong: NOT the reference: the same
    objective with the oracle's analytic sparse Jacobian through least_squares(jac=..., tr_solver='lsmr') at the cfg1 size."""
    from scipy.optimize import least_squares
    import calib_b200  # noqa: F401
    from calib_b200 import scene
    from oracle import ba_oracle as bo
    out = {"host_cpus": os.cpu_count(), "literal": [], "strong": None}
    budget = budget_s
    t_start = time.perf_counter()
    for N in (5, 10, 20, 40):
        if time.perf_counter() - t_start > budget:
            break
        sc = scene.cfg1(seed=300 + N, N=N)
        sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
        C, _, J = ss2d.shape
        x0 = bo.triangulate_with_conf(sp2d, ss2d, sc.K, sc.R_init, sc.t_init.reshape(C, 3, 1), ss2d > 0)
        th0 = bo.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
        args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, 1.0, 1.0)
        t0 = time.perf_counter()
        res = least_squares(bo.objfun, th0, method="trf", ftol=1e-4, args=args)
        dt = time.perf_counter() - t0
        n_obs = int((ss2d > 0).sum())
        out["literal"].append({"frames": N, "n": int(th0.size), "m": int(res.fun.size), "observations": n_obs,
                               "seconds": dt, "iterations": int(res.njev), "nfev": int(res.nfev),
                               "s_per_iteration": dt / res.njev, "obs_per_s_per_iteration": n_obs * res.njev / dt,
                               "status": int(res.status)})
    # strong CPU row at cfg1
    sc = scene.cfg1(seed=0, N=int(os.environ.get("CPU_ROWS_STRONG_N", "1000")))
    sp2d, ss2d, sp3d, ss3d = scene.ba_inputs(sc)
    C, N, J = ss2d.shape
    t0 = time.perf_counter()
    x0 = bo.triangulate_with_conf(sp2d, ss2d, sc.K, sc.R_init, sc.t_init.reshape(C, 3, 1), ss2d > 0)
    t_tri = time.perf_counter() - t0
    th0 = bo.to_theta(sc.R_init, sc.t_init, x0.reshape(-1, 3))
    args = (sc.K, sp2d.reshape(C, N * J, 2), ss2d, sp3d, ss3d, bo.OP_BONE, C, N, J, 1.0, 1.0)
    t0 = time.perf_counter()
    res = least_squares(bo.objfun, th0, jac=bo.jacobian, method="trf", tr_solver="lsmr", ftol=1e-4, args=args,
                        max_nfev=int(os.environ.get("CPU_ROWS_STRONG_NFEV", "12")))
    dt = time.perf_counter() - t0
    n_obs = int((ss2d > 0).sum())
    out["strong"] = {"frames": N, "n": int(th0.size), "observations": n_obs, "seconds": dt, "iterations": int(res.njev),
                     "nfev": int(res.nfev), "s_per_iteration": dt / res.njev,
                     "obs_per_s_per_iteration": n_obs * res.njev / dt, "status": int(res.status),
                     "triangulation_s": t_tri,
                     "note": "analytic sparse Jacobian + LSMR trust-region solver; not what the reference runs"}
    return out



def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="cfg4shard", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "cpu_rows"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-cpu-rows", dest="no_cpu_rows", action="store_true", help="reference arm: skip the executed CPU rows")
    ap.add_argument("--no-other", dest="no_other", action="store_true", help="skip the short cfg1 / cfg3 runs of the one-GPU line")
    ap.add_argument("--batch", type=int, default=512, help="problems per GPU of the sweep leg (0 = skip; 512 x 8 GPUs = BASELINE configs[4])")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "cpu_rows":
        run_cpu_rows(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
