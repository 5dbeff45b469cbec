#!/usr/bin/env python
"""bench.py -- IK ticks/s of the batched IntegrationBasedIK tick (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one control tick (FK + Jacobians -> task assembly -> QP -> ForwardEuler) over one
batch of synthetic robots: the headline workload C3 of BASELINE.json (ergoCub-23, CoM + 2xSE3 feet
+ JointLimits + JointVelocityLimits hard constraints, SO3 chest + JointTracking costs), 65,536
instances PER GPU (weak scaling: the batch is sharded by instance index, no collective on the data
path).  Every step restores the same initial state on the device and solves cold (no warm start),
so all K steps do identical work.

`value`  : device-resident inputs, CUDA-event time of the K steps, max over ranks.
`e2e`    : the same step through the public host-buffer API: pinned host -> device copies of the
           state and set-points, the tick, and the device -> host read of nu + status, all timed.
`--impl reference`: the reference's CPU implementation cannot be built here (no Eigen / iDynTree /
           OSQP / manif in the image), so this arm times the CPU oracle -- the C restatement of the
           reference path -- on all host cores, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

PER_GPU_INSTANCES = 65536
METRIC = "IK ticks/sec (robot-instances x ticks) at 65,536 ergoCub batch on 1/2/4/8 B200"  # BASELINE.json: metric
CONFIG_ID = 3


# ------------------------------------------------------------------ helpers
def algorithmic_flops_per_tick(n: int, me: int, K: int, iters_total: np.ndarray, fast_path: bool, admm_iters_per_round: int = 10,
                               polish_iters: int = 4) -> float:
    """FP64 flops of the algorithm as designed (symmetry / triangular structure exploited, 1 FMA = 2 flops),
    summed over the batch.  DESIGN.md section 4 derives every term.  `iters_total` is the per-instance count of
    N mat-vecs the kernel reports (ADMM + polish iterations)."""
    nv = n + 6
    kin = 4000.0 + 60.0 * n                       # FK, frame Jacobians, CoM subtree pass, controllers
    it = 2.0 * nv * nv + 12.0 * nv                # N r + x_b and the lane-local updates
    if fast_path:
        # G = E + B D^-1 B' (K^2 nv), Cholesky (K^3/3), U (K^2 nv), N = D^-1 - U U' (nv^2 K), x_b
        setup = 2.0 * K * K * nv + K ** 3 / 3.0 + nv * nv * K + 2.0 * K * nv + K * K
        per = kin + 6.0 * n + setup + iters_total * it
    else:
        asm = 2.0 * (K - me) * nv * nv + 6.0 * n  # dense cost rows into H
        setup = nv ** 3 + 3.0 * me * nv * nv + 2.0 * me * me * nv + me ** 3 / 3.0
        cert = 2.0 * nv * nv + 4.0 * me * nv + me * me
        rounds = np.maximum(1, np.ceil(iters_total / max(1, admm_iters_per_round)))
        per = kin + asm + rounds * (2.0 * setup + polish_iters * it + cert) + iters_total * it
    return float(np.sum(per))


def algorithmic_bytes_per_tick(n: int, sp_doubles: int) -> float:
    """HBM bytes one instance-tick must move: state in, set-points in, nu + state + status out."""
    nv = n + 6
    return 8.0 * ((7 + n) + sp_doubles + nv + (7 + n)) + 8.0


class ClockSampler(threading.Thread):
    """Samples nvidia-smi SM clocks / throttle reasons while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.1)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = sorted(float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit())
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        reasons = set()
        for s in self.samples:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def cpu_oracle_rate(workload, inputs, sample: int, threads: int, repeats: int = 1, osqp_defaults: bool = False):
    """ticks/s of the CPU oracle on `sample` instances, cold start.  Default: the accuracy the GPU run is held to (eps 1e-7 +
    polish); osqp_defaults=True: the settings the unmodified reference runs with (OSQP 0.6.3 defaults,
    eps 1e-3, rho 0.1, sigma 1e-6, no polish -- QPInverseKinematics.cpp:339 only sets the verbosity)."""
    from oracle.oracle import Oracle
    orc = Oracle(workload.model, workload.tasks)
    st = orc.settings(parity=not osqp_defaults)
    sub = inputs.slice(0, sample)
    best = 0.0
    for _ in range(repeats):
        p, q, s = sub.base_pos.copy(), sub.base_quat.copy(), sub.s.copy()
        t0 = time.perf_counter()
        _, status, _ = orc.tick_batch(p, q, s, sub.setpoints, st=st, dt_integrate=workload.dt, nticks=1, nthreads=threads)
        dt = time.perf_counter() - t0
        best = max(best, sample / dt)
    return best, int((status != 0).sum())


# ------------------------------------------------------------------ reference arm
def run_reference(args, rank: int, world: int) -> None:
    if rank != 0:
        return
    import blf_b200  # noqa: F401
    from blf_b200 import workloads as W
    w = W.config(CONFIG_ID)
    cores = os.cpu_count() or 1
    sample = 512 * cores
    inputs = W.synth_inputs(w, sample, CONFIG_ID)
    warm_rate = 0.0
    for _ in range(max(1, args.warmup)):
        warm_rate = cpu_oracle_rate(w, inputs, min(sample, 64 * cores), cores)[0]
    steps = max(1, args.steps)
    # every step is a bounded sample of the workload; the sample shrinks if K steps of it would not end within ~2 minutes
    budget_s = 120.0
    if warm_rate > 0 and steps * sample / warm_rate > budget_s:
        sample = max(16 * cores, int(budget_s * warm_rate / steps) // cores * cores)
        inputs = inputs.slice(0, sample)
    t0 = time.perf_counter()
    rates = [cpu_oracle_rate(w, inputs, sample, cores)[0] for _ in range(steps)]
    wall = time.perf_counter() - t0
    value = sample * steps / sum(sample / r for r in rates)
    stock_rate = cpu_oracle_rate(w, inputs, sample, cores, osqp_defaults=True)[0]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "ticks/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": max(1, args.warmup), "ms_per_step": 1e3 * wall / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"C3 ergoCub-23 IK tick (m=107 reference rows), bounded sample of {sample} instances per step",
                   "instances_per_step": sample, "dt": w.dt},
        "cpu_baseline": {"value": value, "unit": "ticks/s", "cores": cores, "kind": "port",
                         "sample": f"{sample} instances x {steps} steps, CPU oracle (C restatement of the reference path; "
                                   "the reference binary cannot be built here), OSQP-equivalent ADMM eps 1e-7 + polish, cold start",
                         "osqp_defaults_value": stock_rate,
                         "osqp_defaults_note": "same sample at the settings the unmodified reference runs with (eps 1e-3, no polish); "
                                               "looser than the accuracy the GPU arm is held to"},
        "e2e": {"value": value, "unit": "ticks/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ the reference-named C++ call
def host_api_leg(w, inputs, devices, steps, warmup, compact, nu_ref=None, status_ref=None, warm_start=False):
    """C3 through QPInverseKinematics::advance() of the C++ host layer (bipedal-locomotion-framework_b200/host/bench/host_bench.cpp):
    a single process drives all `devices`; the batch of `inputs` is what it is given (the caller sizes it)."""
    import tempfile
    host = os.path.join(ROOT, "bipedal-locomotion-framework_b200", "host")
    exe = os.path.join(host, "bench", "host_bench")
    model_file = os.path.join(ROOT, "tests", "golden", "ergocub23.model.txt")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", host, "-s"])
    sp = inputs.setpoints
    order = [0, 1, 2, 5, 6]   # CoM, right foot, left foot, chest, postural (workloads.config(3) task indices)
    with tempfile.TemporaryDirectory() as td:
        fin, fout = os.path.join(td, "inputs.bin"), os.path.join(td, "nu.bin")
        with open(fin, "wb") as f:
            np.array([inputs.base_pos.shape[0], w.model.n], dtype=np.int64).tofile(f)
            for a in [inputs.base_pos, inputs.base_quat, inputs.s] + [sp[t] for t in order]:
                np.ascontiguousarray(a, dtype=np.float64).tofile(f)
        r = subprocess.run([exe, model_file, fin, str(steps), str(warmup), ",".join(str(d) for d in devices), "1" if compact else "0", "1" if warm_start else "0", fout],
                           capture_output=True, text=True, timeout=1200)
        if r.returncode != 0:
            return {"error": f"host_bench exited with {r.returncode}: {(r.stdout + r.stderr)[-400:]}"}
        out = json.loads(r.stdout.strip().splitlines()[-1])
        if nu_ref is not None:
            N, nv = nu_ref.shape
            raw = np.fromfile(fout, dtype=np.uint8)
            nu = raw[:N * nv * 8].view(np.float64).reshape(N, nv)
            st = raw[N * nv * 8:].view(np.int32)
            ok = (st == 0) & (status_ref == 0)
            # the C++ layer adds the tasks in the order of its parameter file, so the rows of the QP are permuted: same optimum,
            # different rounding
            out["status_matches_device"] = bool(np.array_equal(st, status_ref))
            out["max_abs_dnu_vs_device"] = float(np.abs(nu - nu_ref)[ok].max()) if ok.any() else None
    return out


# ------------------------------------------------------------------ our arm
def workload_plan(args, world: int):
    """(config id, instances per GPU, scaling) of this run.  Default: C3, 65,536 instances PER GPU (weak), the line the driver
    records.  BASELINE.json's other shapes: C2 4,096 instances on one GPU; C3 strong: 65,536 in total; C4 262,144 over 8 GPUs =
    32,768 per GPU; C5 65,536 over 8 GPUs (8,192 per GPU), 1,000-tick closed-loop launches."""
    cid = args.config
    scaling = args.scaling or {2: "strong", 3: "weak", 4: "weak", 5: "strong"}[cid]
    base = {2: 4096, 3: 65536, 4: 32768, 5: 65536}[cid]
    if args.instances:
        n_local = args.instances
    elif scaling == "weak":
        n_local = base
    else:
        n_local = max(1, base // world)
    return cid, n_local, scaling


def run_ours(args, rank: int, local_rank: int, world: int) -> None:
    import torch
    import torch.distributed as dist
    import blf_b200  # noqa: F401
    from blf_b200 import capi, workloads as W, tasks as T

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    cid, n_local, scaling = workload_plan(args, world)
    w = W.config(cid)
    n, nv = w.model.n, w.model.n + 6
    n_global = n_local * world
    lo, hi = W.shard_range(n_global, rank, world)
    # global synthetic batch, sliced by instance range: data of instance i are independent of `world`
    if args.inputs == "survey":   # SURVEY.md 8d as written: joints over the inner 80 % of their range, +-5 cm / +-0.1 rad, CoM +-3 cm
        inputs = W.synth_inputs_survey(w, n_global, cid).slice(lo, hi)
    else:
        inputs = W.synth_inputs(w, n_global, cid).slice(lo, hi)
    rollout = cid == 5
    nticks = 1000 if rollout else 1

    model = capi.Model(w.model)
    prob = capi.Problem(model, w.tasks, n_local, device=local_rank)
    stream = torch.cuda.ExternalStream(prob.stream, device=local_rank)
    dev = torch.device("cuda", local_rank)
    d_state = [torch.from_numpy(a).to(dev) for a in (inputs.base_pos, inputs.base_quat, inputs.s)]
    h_nu = torch.empty((n_local, nv), dtype=torch.float64).pin_memory()
    h_status = torch.empty(n_local, dtype=torch.int32).pin_memory()
    sp_doubles = sum(0 if sp is None else sp.shape[1] for sp in inputs.setpoints)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    prob.set_inputs(inputs)
    if rollout:
        # C5: "random SE3 target trajectories" generated on the device (sinusoids, period 2 s, random phase), closed loop,
        # warm-started, nothing crosses the host link inside a launch.  SURVEY 8d prescribes amplitudes of 0.05 m / 0.1 rad
        # (--inputs survey); with both feet and the CoM as hard constraints and +-2 rad/s velocity limits that makes a fifth of
        # the robots infeasible from the first ticks on (each then burns the iteration budget every tick), so the default
        # moves the feet by 0.02 m / 0.04 rad (99.5 % of the robots solved at every one of the 1,000 ticks)
        amp_l, amp_a = (0.05, 0.1) if args.inputs == "survey" else (0.02, 0.04)
        rng = np.random.Generator(np.random.Philox(key=W.SEED + 50))
        phases = rng.uniform(0.0, 2.0 * np.pi, size=(n_global, len(w.tasks)))[lo:hi]
        for t, task in enumerate(w.tasks):
            if task.type == T.TASK_SE3:
                prob.set_setpoint_trajectory(t, [amp_l] * 3 + [amp_a] * 3, 2.0, np.ascontiguousarray(phases[:, t]))
    prob.sync()

    def device_step():
        prob.set_time(0.0)
        prob.set_state(*d_state)                       # device -> device restore of the initial state
        prob.tick(nticks, w.dt, integrate=True, warm_start=rollout)

    # Host records.  The set-points of the open-loop configs carry no feed-forward velocities ("feed-forward twist 0
    # (open loop)"), which is the reference's setSetPoint(pose) with its default zero velocity: the records then hold the pose
    # part only (BLFIK_TICK_NO_FEEDFORWARD; pack_inputs refuses if a dropped entry is non-zero).  A set-point that is the same
    # vector for every robot -- the postural JointTrackingTask target, q_nominal -- is handed over once
    # (blfik_set_setpoint_shared) and is not part of the records at all.
    compact = not args.full_records
    shared = []
    if not args.no_shared and not rollout:
        for t, sp in enumerate(inputs.setpoints):
            if sp is not None and w.tasks[t].type == T.TASK_JOINT_TRACKING and np.all(sp == sp[0]):
                shared.append(t)
    e2e_prob = capi.Problem(model, w.tasks, n_local, device=local_rank) if shared else prob
    if shared:
        e2e_prob.set_inputs(inputs)
        for t in shared:
            e2e_prob.set_setpoint_shared(t, inputs.setpoints[t][0])
    if compact:
        try:
            packed_np = e2e_prob.pack_inputs(inputs, compact=True)
        except ValueError:
            compact = False
    if not compact:
        packed_np = e2e_prob.pack_inputs(inputs)
    h_packed = torch.from_numpy(packed_np).pin_memory()

    def e2e_step():
        if rollout:
            # the closed-loop rollout has no per-tick host traffic by design; end to end = initial state in (pinned host ->
            # device), the 1,000-tick launch, final state + last nu + status out
            prob.set_time(0.0)
            prob.set_state(inputs.base_pos, inputs.base_quat, inputs.s)
            prob.tick(nticks, w.dt, integrate=True, warm_start=True)
            prob.get_state()
            prob.get_output(h_nu)
            prob.get_status()
        else:
            # the reference-facing call with HOST buffers: one packed record per robot (state + set-points) in, nu + status out
            e2e_prob.tick_packed(h_packed, h_nu, h_status, dt=w.dt, integrate=True, warm_start=False, nchunks=args.chunks, compact=compact)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps, st=stream, queued=False):
        """Sum of the K step durations, each between its own pair of CUDA events on the stream the step runs on; the L2 flush
        between steps is outside the events.  queued=True (device-resident step): flush, events and steps are all enqueued on
        that one stream without a host synchronisation in between, so the device time of a step is measured, not the host's
        call latency in front of an idle GPU.  queued=False (host-buffer step): the call is synchronous by design, every step is
        timed from a quiet device."""
        evs = []
        for _ in range(steps):
            with torch.cuda.stream(st):
                flush.fill_(1)                        # evict L2 (126 MB) between steps, untimed
            if not queued:
                torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            step_fn()
            e1.record(st)
            if not queued:
                e1.synchronize()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    steps = args.steps if not rollout else max(1, min(args.steps, 10))
    for _ in range(max(3, args.warmup) if not rollout else 1):
        device_step()
    prob.sync()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = prob.launch_count
    barrier()
    ms_dev = timed(device_step, steps, queued=True)
    barrier()
    launches = prob.launch_count - launches0
    status, iters = prob.get_status()
    unsolved = int((status != 0).sum())
    nu_dev = prob.get_output()

    # per-kernel durations of the two-launch path, CUDA events on the launching stream (a separate pass: the events sit
    # between the kernels)
    k_asm = k_solve = None
    if prob.kernel_path == 4 and not rollout:
        prob.kernel_timing(True)
        ta, tb = [], []
        for _ in range(min(steps, 20)):
            flush.fill_(1)
            torch.cuda.synchronize()
            device_step()
            a_ms, b_ms = prob.kernel_times()
            ta.append(a_ms), tb.append(b_ms)
        prob.kernel_timing(False)
        k_asm, k_solve = float(np.mean(ta)), float(np.mean(tb))

    e2e_stream = torch.cuda.ExternalStream(e2e_prob.stream, device=local_rank)
    for _ in range(2 if not rollout else 1):
        e2e_step()
    barrier()
    ms_e2e = timed(e2e_step, steps, e2e_stream)
    barrier()
    clocks = sampler.stop() if sampler else None
    # the numbers the end-to-end path delivered, against the device-resident step's (same inputs, same cold start)
    e2e_ok = bool(np.array_equal(h_status.numpy(), status)) and bool(np.array_equal(h_nu.numpy(), nu_dev))

    t = torch.tensor([ms_dev, ms_e2e, float(unsolved), 0.0 if e2e_ok else 1.0], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms_dev, ms_e2e, unsolved, e2e_ok = float(tmax[0]), float(tmax[1]), int(tsum[2]), float(tsum[3]) == 0.0
    ticks_total = n_global * steps * nticks

    if rank == 0:
        fp64_peak = capi.measure_fp64_peak(local_rank, 300.0)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
        K = sum(tk.rows(n) for tk in w.tasks if tk.type in (0, 1, 2, 6))
        kp = prob.kernel_path
        fast = kp in (2, 4)
        it64 = iters.astype(np.float64) / nticks
        flops_tick = algorithmic_flops_per_tick(n, prob.me, K, it64, fast, prob_settings["admm_iters"], prob_settings["polish_iters"])
        flops_kin = (4000.0 + 60.0 * n + 6.0 * n) * n_local
        step_ms = ms_dev / steps
        if k_solve:   # the dominant kernel of the two-launch path: K3 + K4; its flops are the tick's minus kinematics / assembly
            kernel_name, kernel_ms, flops_launch = f"blfik_solve_kernel<{nv},{18 if K <= 18 else 32}>", k_solve, flops_tick - flops_kin
        else:
            kernel_name = {1: "blfik_tick_kernel", 2: "blfik_tick_fast_kernel", 3: "blfik_tick_eq_kernel", 4: "blfik_solve_kernel"}[kp]
            kernel_ms, flops_launch = step_ms / nticks, flops_tick
        achieved_tf = flops_launch / (kernel_ms * 1e-3) * 1e-12
        bytes_tick = algorithmic_bytes_per_tick(n, sp_doubles) * n_local
        cores = os.cpu_count() or 1
        cpu_sample = min(4096 * cores, n_local)   # 65,536 instances on a 16-core box: about 20 s of CPU work at the GPU run's accuracy
        cpu_rate, cpu_bad = cpu_oracle_rate(w, inputs, cpu_sample, cores)
        cpu_stock = cpu_oracle_rate(w, inputs, cpu_sample, cores, osqp_defaults=True)[0]
        # issue roofline: the warp-instructions of this step's kernels against what the schedulers can issue (4 per clock and SM)
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
        line = {
            "metric": METRIC, "value": ticks_total / (ms_dev * 1e-3), "unit": "ticks/s", "n_gpus": world,
            "steps": steps, "warmup": max(3, args.warmup), "ms_per_step": ms_dev / steps,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"C{cid}: {w.description}" + (", 1,000-tick closed-loop launches (ForwardEuler on the device, "
                                   "sinusoidal SE3 targets generated on the device, warm start)" if rollout else ", then ForwardEuler"),
                       "config_id": cid, "instances_per_gpu": n_local, "instances_total": n_global, "ticks_per_step": nticks, "dt": w.dt,
                       "inputs": ("SURVEY 8d distribution as written (joints over the inner 80 % of their range, +-5 cm / +-0.1 rad, CoM "
                                  "+-3 cm)" if args.inputs == "survey" else "workloads.synth_inputs (nominal +-0.25 rad, +-3 cm / +-0.06 rad, CoM +-2 cm)"),
                       "parallelism": f"instance-sharded x{world}, no collective", "warm_start": rollout,
                       "solver": prob_settings, "l2": "256 MiB write between steps (untimed) flushes the 126 MB L2",
                       "unsolved_instances": unsolved, "kernel_path": kp,
                       "host_records": ("no per-tick host traffic (closed loop on the device)" if rollout else
                                        (f"compact ({h_packed.shape[1] * 8} B per robot): state + pose part of every set-point" if compact
                                         else f"full ({h_packed.shape[1] * 8} B per robot)") +
                                        (f"; batch-shared set-points sent once, not per step: tasks {shared}" if shared else "")),
                       "host_tick": ("initial state in, one 1,000-tick launch, final state + nu + status out" if rollout else
                                     f"blfik_tick_packed, {args.chunks} pipeline chunks: record copies (copy engine), assemble + solve "
                                     "kernels per chunk on alternating streams, results written to pinned host memory by the kernel")},
            "e2e": {"value": ticks_total / (ms_e2e * 1e-3), "unit": "ticks/s",
                    "h2d_bytes_per_step": int(h_packed.numel() * 8) if not rollout else int(n_local * (7 + n) * 8),
                    "d2h_bytes_per_step": int(n_local * (8 * nv + 4)) + (int(n_local * (7 + n) * 8) if rollout else 0)},
            "e2e_matches_device": e2e_ok if not rollout else None,
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp64", "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": achieved_tf / fp64_peak if fp64_peak else None,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one launch of that kernel (profiles/r2_*_ncu_summary.txt)
                         "traffic": NCU_TRAFFIC.get((kernel_name.split("<")[0], cid, n_local)),
                         "peak_source": "measured live by blfik_measure_fp64_peak (dependency-free DFMA loop); "
                                        "MEASURED_PEAKS.json has no FP64 figure",
                         "kernel": kernel_name, "kernel_ms": kernel_ms, "kernel_timing": "CUDA events on the launching stream around the kernel"
                                   if k_solve else "CUDA events around the step", "flops_per_launch": flops_launch,
                         "kernels_ms": {"blfik_assemble_kernel": k_asm, "blfik_solve_kernel": k_solve} if k_solve else None,
                         "mean_matvecs_per_tick": float(it64.mean()),
                         "whole_step": {"achieved": flops_tick / (step_ms / nticks * 1e-3) * 1e-12,
                                        "frac": flops_tick / (step_ms / nticks * 1e-3) * 1e-12 / fp64_peak if fp64_peak else None},
                         "issue": {"note": "warp-instructions of one step (ncu smsp__inst_executed.sum of both kernels, profiles/) over the "
                                           "issue rate 4 x SMs x SM clock", "warp_instructions": NCU_INSTR.get((cid, n_local)),
                                   "floor_ms": (NCU_INSTR[(cid, n_local)] / (4 * 148 * sm_mhz * 1e6) * 1e3) if (cid, n_local) in NCU_INSTR else None},
                         "hbm": {"achieved": bytes_tick / (step_ms / nticks * 1e-3) * 1e-9, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": bytes_tick / (step_ms / nticks * 1e-3) * 1e-9 / hbm_peak, "peak_source": hbm_src,
                                 "bytes_per_launch": bytes_tick}},
            "cpu_baseline": {"value": cpu_rate, "unit": "ticks/s", "cores": cores, "kind": "port",
                             "sample": f"{cpu_sample} instances x 1 tick of the same workload, CPU oracle "
                                       f"(eps 1e-7 + polish, cold start), {cpu_bad} unsolved",
                             "osqp_defaults_value": cpu_stock,
                             "osqp_defaults_note": "same sample at the settings the unmodified reference runs with (OSQP defaults: eps 1e-3, "
                                                   "no polish), looser than the accuracy this run is held to"},
            "clocks": clocks,
        }
        if cid == 3 and world == 1 and not args.no_host_leg:
            # the same step through the reference-named C++ call (one process, one device here; `--via host --gpus N` for more)
            try:
                line["e2e_host_api"] = host_api_leg(w, inputs, [local_rank], min(steps, 20), 3, compact, nu_dev, status)
            except Exception as e:   # the leg is an extra: never let it take the line down
                line["e2e_host_api"] = {"error": repr(e)}
        print(json.dumps(line), flush=True)
    if e2e_prob is not prob:
        e2e_prob.close()
    prob.close()
    if world > 1:
        dist.destroy_process_group()


# ncu captures of the shipped build (profiles/): DRAM bytes of one launch and warp-instructions of one step, keyed by workload
NCU_TRAFFIC = {   # profiles/r2c_split_kernels_ncu_summary.txt, profiles/r2_eq_kernel_v3_ncu_summary.txt
    ("blfik_solve_kernel", 3, 65536): 387.19e6 + 66.61e6,
    ("blfik_assemble_kernel", 3, 65536): 185.93e6 + 474.26e6,
    ("blfik_tick_eq_kernel", 2, 4096): 5.10e6 + 0.03e6,
}
NCU_INSTR = {(3, 65536): 70455143 + 302609305}   # smsp__inst_executed.sum: assemble + solve

prob_settings = {"rho": 1.0, "sigma": 1e-2, "alpha": 1.6, "eps": 1e-7, "admm_iters": 10, "max_rounds": 8,
                 "polish_iters": 4, "eq_iters": 5}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--instances", type=int, default=0, help="instances per GPU (default: what the configuration prescribes)")
    ap.add_argument("--config", type=int, default=CONFIG_ID, choices=[2, 3, 4, 5], help="BASELINE.json configuration (default 3, the headline)")
    ap.add_argument("--scaling", default=None, choices=["weak", "strong"], help="weak: fixed instances per GPU; strong: fixed total (defaults per configuration)")
    ap.add_argument("--inputs", default="default", choices=["default", "survey"], help="synthetic input distribution (survey: SURVEY.md 8d as written)")
    ap.add_argument("--no-host-leg", action="store_true", help="skip the extra leg through the C++ QPInverseKinematics::advance()")
    ap.add_argument("--via", default="capi", choices=["capi", "host"], help="host: time ONLY the C++ host-layer call, one process driving --gpus devices (weak: 65,536 instances per GPU)")
    ap.add_argument("--no-shared", action="store_true", help="host-buffer (e2e) tick: send the batch-shared set-points with every record")
    ap.add_argument("--full-records", action="store_true", help="host-buffer (e2e) tick: send full records (with the zero feed-forward terms) instead of compact ones")
    ap.add_argument("--chunks", type=int, default=8,
                    help="host-buffer (e2e) tick: k >= 1 = pipelined (pinned buffers: one persistent launch that consumes the records while the copy engine streams them in; otherwise k chunk launches); 0 = zero-copy, the kernel reads the records over PCIe itself")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.via == "host":
        if rank == 0:
            import blf_b200  # noqa: F401
            from blf_b200 import workloads as W
            w = W.config(3)
            per = args.instances or PER_GPU_INSTANCES
            inputs = W.synth_inputs(w, per * args.gpus, 3)
            out = host_api_leg(w, inputs, list(range(args.gpus)), args.steps, max(3, args.warmup), not args.full_records)
            out.update({"metric": METRIC, "unit": "ticks/s", "value": out.get("ticks_per_s"), "n_gpus": args.gpus, "scaling": "weak",
                        "config": {"workload": "C3 through QPInverseKinematics::advance() (C++ host layer), one process, one host thread and one "
                                               "blfik_problem per GPU, contiguous instance blocks", "instances_per_gpu": per}})
            print(json.dumps(out), flush=True)
        return
    if world == 1 and args.gpus > 1:
        # convenience: relaunch under torchrun, one rank per GPU
        cmd = ([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__),
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup), "--instances", str(args.instances),
               "--chunks", str(args.chunks), "--config", str(args.config), "--inputs", args.inputs] +
               (["--scaling", args.scaling] if args.scaling else []) + (["--full-records"] if args.full_records else []) +
               (["--no-shared"] if args.no_shared else []))
        raise SystemExit(subprocess.call(cmd))
    run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
