#!/usr/bin/env python
"""bench.py -- headline benchmark of the sampling-MPC hot path on B200 (contract: see DESIGN.md "Measurement").

Workload (BASELINE.json configs[1], "C2"): Go2, bound gait, MPPI, cubic-spline parameters (P = 96), per-sample
step-frequency adaptation, 10,000 rollouts over a 12-step horizon.  One "step" = one MPC solve of that robot.

  value        rollouts/s with the inputs resident in HBM (qpmpc_solve_batched on device tensors), device-timed
               per step with CUDA events on the handle's stream; L2 is flushed between timed steps.
  e2e          the same metric through the reference-facing call qpmpc_solve with HOST buffers, wall clock per call:
               the inputs (1176 B) go host -> device inside the launch (kernel parameter block), the one kernel runs,
               and its finalize step stores the results device -> host into mapped pinned memory as {word, sequence}
               pairs the host waits on (QPMPC_ZEROCOPY=0: H2D copy + kernel + D2H copy + stream sync); p50 reported.
  roofline     k_rollout against the FP32 CUDA-core peak measured in this run (no tensor cores: nothing on this path
               is a dense contraction); algorithmic FLOPs per rollout from SURVEY.md 8(d).
  cpu_baseline the CPU oracle port (-O3, OpenMP over rollouts, all host threads) on the same workload.
  batched      BASELINE.json configs[4] ("C5", 8192 robots x 1024 rollouts, MPPI zero-order), sharded by robot over
               the ranks with no collective on the control path.

`--impl reference` times the reference's CPU implementation of the path.  jax is not installable here, so that arm
is the oracle port of the reference's jax code (kind "port"), all host threads.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GO2 = dict(mass=15.019, inertia=np.array([[1.58460467e-01, 1.21660000e-04, -1.55444692e-02],
                                          [1.21660000e-04, 4.68645637e-01, -3.12000000e-05],
                                          [-1.55444692e-02, -3.12000000e-05, 5.24474661e-01]]))
ALIENGO = dict(mass=24.637, inertia=np.array([[0.2310941359705289, -0.0014987128245817424, -0.021400468992761768],
                                              [-0.0014987128245817424, 1.4485084687476608, 0.0004641447134275615],
                                              [-0.021400468992761768, 0.0004641447134275615, 1.503217877350808]]))
N_ROLLOUTS = 10000
FLOP_PER_ROLLOUT = {"zero_order": 4380, "linear_spline": 4956, "cubic_spline": 7260, "cubic_spline_ga": 7550}
WORKLOAD = "C2: go2 bound, MPPI, cubic_spline (P=96), per-sample step-frequency adaptation, 10000 rollouts x horizon 12"


def synth_problem(rng, h, hip, gait=([0.5, 0.5, 0.0, 0.0], 1.8, 0.65), horizon=12, dt=0.02):
    """SURVEY 8(d) state / reference distribution and a host gait-clock contact sequence."""
    hx, hy = hip
    st = np.zeros(24)
    st[0:3] = [0.0, 0.0, h + rng.normal(0, 0.02)]
    st[3:6], st[6:9], st[9:12] = rng.normal(0, 0.3, 3), rng.normal(0, 0.05, 3), rng.normal(0, 0.3, 3)
    hips = np.array([[hx, hy, 0], [hx, -hy, 0], [-hx, hy, 0], [-hx, -hy, 0]], float)
    st[12:24] = (hips + np.c_[rng.normal(0, 0.03, (4, 2)), np.zeros(4)]).ravel()
    ref = np.zeros(24)
    ref[2], ref[3], ref[12:24] = h, rng.uniform(-0.5, 0.5), st[12:24]
    offs, freq, duty = gait
    t = (np.array(offs) + rng.uniform()) % 1.0
    timing = t.copy()
    seq = np.zeros((4, horizon))
    for n in range(horizon):
        seq[:, n] = (t < duty).astype(float)
        t = (t + dt * freq) % 1.0
    return st.astype(np.float32), ref.astype(np.float32), seq.astype(np.float32), timing.astype(np.float32), freq


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------------------------------- reference arm (CPU)
def cpu_solve_loop(steps, warmup, seed=0):
    """The CPU port of compute_control_mppi on the C2 workload: noise generation + 10k rollouts + update per step."""
    from oracle import srbd_oracle as so
    so.set_threads(os.cpu_count() or 1)  # all host threads, also under torchrun (which exports OMP_NUM_THREADS=1)
    cfg = so.make_cfg(parametrization="cubic_spline", gait_adaptive=True, **GO2)
    rng = np.random.default_rng(seed)
    st, ref, seq, timing, nominal = synth_problem(rng, 0.27, (0.19, 0.13))
    P = so.num_params(cfg)
    best = np.zeros(P, np.float32)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        eps = so.fill_noise(N_ROLLOUTS, P, 3.0, it, fast=True)
        freqs = np.asarray([1.4, 2.0, 2.4], np.float32)[(np.arange(N_ROLLOUTS) * 7 + it) % 3]
        r = so.solve(cfg, "mppi", st, ref, seq, best, eps, timing=timing, freqs=freqs, fast=True)
        best = r["best"]
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    return np.array(times), so.max_threads(True)


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    steps = max(args.steps, 1)
    times, cores = cpu_solve_loop(steps, args.warmup)
    ms = float(times.mean() * 1e3)
    value = N_ROLLOUTS / times.mean()
    sample = f"{steps} full solves of the C2 workload (noise + 10000 rollouts + MPPI update each), OpenMP over rollouts"
    line = {"impl": "reference", "metric": "rollouts/sec at 10k samples (one MPPI solve)", "value": value, "unit": "rollouts/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "p50_latency_ms": float(np.median(times) * 1e3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "note": "reference jax unavailable on this box; CPU port of the reference math (oracle/srbd_oracle.c, -O3 -fopenmp)"},
            "cpu_baseline": {"value": value, "unit": "rollouts/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "rollouts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------- clocks
class Clocks:
    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self.h = [], set(), None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def sample(self):
        if self.h is None:
            return
        nv = self.nv
        self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
        try:
            bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:  # noqa: BLE001
            bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake_slowdown"}
        for bit, name in names.items():
            if bits & bit:
                self.reasons.add(name)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------- CUDA arm
def run_cuda(args):
    import torch
    from quadruped_pympc_tamols_b200 import _cabi
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    K, W = max(args.steps, 1), max(args.warmup, 3)
    rng = np.random.default_rng(1000 + rank)  # each rank = an independent robot instance (batched_simulations.py style)
    st, ref, seq, timing, nominal = synth_problem(rng, 0.27, (0.19, 0.13))
    h = _cabi.Handle(_cabi.make_config(sampling_method="mppi", parametrization="cubic_spline", gait_adaptive=True,
                                       num_parallel_computations=N_ROLLOUTS, device=local_rank, **GO2))
    P = h.P
    stream = torch.cuda.ExternalStream(h.stream, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    clocks = Clocks(local_rank)

    # ---- device-resident arm
    t_ = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=dev)  # noqa: E731
    d_state, d_ref, d_seq, d_timing = t_(st[None]), t_(ref[None]), t_(seq[None]), t_(timing[None])
    d_nom, d_osw = t_(np.array([nominal])), t_(np.array([1]), torch.int32)
    d_best = [torch.zeros((1, P), device=dev), torch.zeros((1, P), device=dev)]
    d_res = torch.zeros((1, _cabi.RESULT_FLOATS), device=dev)
    launches0 = h.launch_count

    def device_step(it):
        b = _cabi.QpmpcBatch(num_robots=1, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(),
                             d_contact_sequence=d_seq.data_ptr(), d_best_in=d_best[it & 1].data_ptr(),
                             d_timing=d_timing.data_ptr(), d_nominal_freq=d_nom.data_ptr(),
                             d_optimize_swing=d_osw.data_ptr(), d_result=d_res.data_ptr(),
                             d_best_out=d_best[(it + 1) & 1].data_ptr())
        h.solve_batched(b, 42, it)

    with torch.cuda.stream(stream):
        for it in range(W):
            device_step(it)
        barrier()
        launches0 = h.launch_count
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        clocks.sample()
        for it in range(K):
            if not os.environ.get('QPMPC_BENCH_NOFLUSH'):
                flush.fill_(it & 0xFF)  # untimed L2 flush between timed steps
            ev[it][0].record(stream)
            device_step(W + it)
            ev[it][1].record(stream)
            if it % max(1, K // 8) == 0:
                clocks.sample()
        clocks.sample()
        barrier()
    dev_ms = np.array([a.elapsed_time(b) for a, b in ev])
    launches = h.launch_count - launches0
    total_ms = max_over_ranks(float(dev_ms.sum()))
    value = N_ROLLOUTS * K * world / (total_ms * 1e-3)

    # ---- roofline of the dominant kernel (k_rollout), events around each kernel inside the solve
    h.set_profiling(True)
    roll = []
    with torch.cuda.stream(stream):
        for it in range(min(K, 200)):
            flush.fill_(it & 0xFF)
            device_step(W + K + it)
            roll.append(h.profile_last())
    h.set_profiling(False)
    roll = np.array(roll)
    fp32_peak = _cabi.measure_fp32_peak(local_rank)
    flop = FLOP_PER_ROLLOUT["cubic_spline_ga"] * N_ROLLOUTS
    ach = flop / (roll[:, 0].mean() * 1e-3) / 1e12

    # ---- end to end through the host-buffer C ABI call (inputs down, one kernel, results back up), wall clock
    best = np.zeros(P, np.float32)
    e2e = []
    for it in range(W + K):
        t0 = time.perf_counter()
        r = h.solve(st, ref, seq, best, 42, it, timing=timing, nominal_freq=nominal, optimize_swing=1, want_costs=False)
        dt_ = time.perf_counter() - t0
        best = r["best"]
        if it >= W:
            e2e.append(dt_)
    e2e = np.array(e2e)
    e2e_total = max_over_ranks(float(e2e.sum()))
    e2e_value = N_ROLLOUTS * K * world / e2e_total
    in_bytes = 4 * (24 + 24 + 4 * 12 + 2 * P + 4 + 2)

    # ---- num_sampling_iterations = 3 chained on the device (qpmpc_solve_iterated) against three
    # host-level calls (the upstream interface loop); extra information, not the headline
    def timed(fn, n=K):
        ts = []
        for it in range(W + n):
            t0 = time.perf_counter()
            fn(it)
            if it >= W:
                ts.append(time.perf_counter() - t0)
        return float(np.median(ts) * 1e3)

    def three_calls(it):
        b_ = best
        for k in range(3):
            b_ = h.solve(st, ref, seq, b_, 42, 3 * it + k, timing=timing, nominal_freq=nominal, optimize_swing=1,
                         want_costs=False)["best"]

    # the same solve one level up: Sampling_MPC (gait-adaptive class) driven like srbd_controller_interface.py:131-168
    from quadruped_pympc_tamols_b200._config import make_config as mk_cfg
    from quadruped_pympc_tamols_b200.controllers.sampling.centroidal_nmpc_jax_gait_adaptive import Sampling_MPC
    ctl = Sampling_MPC(config=mk_cfg("go2", sampling_method="mppi", control_parametrization="cubic_spline",
                                     optimize_step_freq=True, num_parallel_computations=N_ROLLOUTS), device_index=local_rank)

    def class_call(it):
        ctl.with_newkey()
        out_ = ctl.jitted_compute_control(st, ref, seq, ctl.best_control_parameters, ctl.master_key, timing, nominal, 1)
        ctl.best_control_parameters = out_[3]

    class_p50 = timed(class_call)
    ctl.close()
    iterated = {"workload": "C2 with num_sampling_iterations = 3 (srbd_controller_interface.py:131-170)",
                "chained_p50_ms": timed(lambda it: h.solve(st, ref, seq, best, 42, 3 * it, timing=timing, nominal_freq=nominal,
                                                           optimize_swing=1, want_costs=False, iterations=3)),
                "three_calls_p50_ms": timed(three_calls)}
    out_bytes = 8 * (_cabi.RESULT_FLOATS + P)  # result record + best_out, each 4-byte word paired with its sequence number

    # ---- batched robots C5 (sharded by robot, no collective): extra information, not the headline
    batched = None
    if not args.no_batched:
        batched = run_batched(torch, _cabi, dev, local_rank, rank, world, barrier, max_over_ranks, fp32_peak)

    tamols = None
    if not args.no_tamols:
        tamols = run_tamols(torch, dev, local_rank, rank, world, barrier, max_over_ranks,
                            with_cpu=(rank == 0 and world == 1 and not args.no_cpu))

    # ---- CPU baseline (rank 0, N=1 only): bounded sample of the same workload
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        nsolve = 4000  # about 10 s of CPU work on the B200 host (2.4 ms per solve with 16 threads)
        times, cores = cpu_solve_loop(nsolve, 3)
        cpu = {"value": N_ROLLOUTS / float(times.mean()), "unit": "rollouts/s", "cores": cores, "kind": "port",
               "sample": f"{nsolve} full C2 solves (noise + 10000 rollouts + MPPI update), {times.sum():.1f} s of CPU work",
               "p50_latency_ms": float(np.median(times) * 1e3)}

    clk = clocks.summary()
    if rank == 0:
        line = {"metric": "rollouts/sec at 10k samples (one MPPI solve)", "value": value, "unit": "rollouts/s", "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": float(total_ms / K), "p50_latency_ms": float(np.median(dev_ms)),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "noise": "in-kernel Philox4x32-10 + Box-Muller", "l2": "flushed (256 MiB write) between timed steps",
                           "multi_gpu": "one independent robot instance per rank, no collective on the control path"},
                "e2e": {"value": e2e_value, "unit": "rollouts/s", "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": out_bytes,
                        "p50_latency_ms": float(np.median(e2e) * 1e3), "p99_latency_ms": float(np.percentile(e2e, 99) * 1e3),
                        "class_call_p50_latency_ms": class_p50,
                        "api": "qpmpc_solve (host buffers; costs[N] left on the device as upstream)",
                        "transfers": "inputs in the kernel parameter block (host -> device with the launch); results stored by the kernel into mapped pinned host memory as {word, sequence} pairs"},
                "gpu_launches": int(launches),
                "roofline": {"bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s", "frac": ach / fp32_peak,
                             "traffic": 77568, "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one launch, profiles/r4c_ncu_raw_c2.csv (bytes)",
                             "kernel": "k_rollout_quad<cubic,gait_adaptive,72> (139 blocks x 288 threads; rollouts + fused finalize, one launch per solve)", "launch_ms": float(roll[:, 0].mean()),
                             "finalize_ms": float(roll[:, 1].mean()), "finalize": "fused into the rollout kernel (last block per robot)", "flop_per_rollout": FLOP_PER_ROLLOUT["cubic_spline_ga"],
                             "peak_source": "FP32 FMA-chain microbenchmark in this run (qpmpc_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 figure",
                             "note": "10k rollouts = 75.5 MFLOP: latency-bound at this size; see batched.roofline for the throughput regime"},
                "cpu_baseline": cpu, "clocks": clk, "iterated": iterated, "batched": batched, "tamols": tamols}
        print(json.dumps(line))
    h.close()
    if world > 1:
        dist.destroy_process_group()


def run_batched(torch, _cabi, dev, local_rank, rank, world, barrier, max_over_ranks, fp32_peak, robots_total=8192, n=1024):
    from quadruped_pympc_tamols_b200.sharding import shard_bounds
    lo, hi = shard_bounds(robots_total, world, rank)  # contiguous block of robots per rank, no collective
    R = hi - lo
    rng = np.random.default_rng(77 + rank)
    h = _cabi.Handle(_cabi.make_config(sampling_method="mppi", parametrization="zero_order", num_parallel_computations=n,
                                       max_robots=R, device=local_rank, **ALIENGO))
    P = h.P
    probs = [synth_problem(rng, 0.33, (0.24, 0.15), gait=([0.5, 1.0, 1.0, 0.5], 1.4, 0.65)) for _ in range(64)]
    idx = rng.integers(0, 64, R)
    t_ = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32, device=dev)  # noqa: E731
    d_state = t_(np.stack([probs[i][0] for i in idx]) + rng.normal(0, 1e-3, (R, 24)).astype(np.float32))
    d_ref, d_seq = t_(np.stack([probs[i][1] for i in idx])), t_(np.stack([probs[i][2] for i in idx]))
    d_best = [torch.zeros((R, P), device=dev), torch.zeros((R, P), device=dev)]
    d_res = torch.zeros((R, _cabi.RESULT_FLOATS), device=dev)
    stream = torch.cuda.ExternalStream(h.stream, device=dev)
    h.set_profiling(True)
    steps, roll = 10, []
    with torch.cuda.stream(stream):
        for it in range(3 + steps):
            b = _cabi.QpmpcBatch(num_robots=R, robot_offset=lo, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(),
                                 d_contact_sequence=d_seq.data_ptr(), d_best_in=d_best[it & 1].data_ptr(),
                                 d_result=d_res.data_ptr(), d_best_out=d_best[(it + 1) & 1].data_ptr())
            if it == 3:
                barrier()
            h.solve_batched(b, 7, it)
            if it >= 3:
                roll.append(h.profile_last())
        barrier()
    roll = np.array(roll)
    step_ms = max_over_ranks(float((roll[:, 0] + roll[:, 1]).mean()))
    roll_ms = max_over_ranks(float(roll[:, 0].mean()))
    ach = FLOP_PER_ROLLOUT["zero_order"] * R * n / (roll_ms * 1e-3) / 1e12
    # ---- the step in front of the rollouts, contact sequences of this rank's robots from R gait clocks
    from quadruped_pympc_tamols_b200.helpers.periodic_gait_generator import PeriodicGaitGenerator
    from oracle import gait_oracle as go  # CPU leg only
    gen = PeriodicGaitGenerator(rng.uniform(0.5, 0.8, R), rng.uniform(1.0, 2.4, R), rng.integers(0, 7, R), 12,
                                num_robots=R, handle=h)
    ph0 = rng.uniform(0, 1, (R, 4))
    gen.set_phase_signal(ph0)
    gait_ms = []
    with torch.cuda.stream(stream):
        for it in range(13):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            gen.compute_contact_sequence([0.02], [12])
            e1.record(stream)
            e1.synchronize()
            if it >= 3:
                gait_ms.append(e0.elapsed_time(e1))
    gait = {"workload": f"contact sequences [R,4,12] of {R} gait clocks per rank (PeriodicGaitGenerator.compute_contact_sequence)",
            "p50_ms": float(np.median(gait_ms)), "sequences_per_s": robots_total / (max_over_ranks(float(np.median(gait_ms))) * 1e-3),
            "algorithmic_bytes": R * (4 * 8 + 4 + 4 * 8 + 8 + 8 + 1 + 4 * 12 * 4),
            "note": "2.3 MB of traffic per launch: launch-latency-bound, HBM time at the measured peak is 0.35 us"}
    if rank == 0 and world == 1:
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 2.0:
            go.contact_sequences(ph0, np.zeros((R, 4), bool), gen.phase_offset, gen.step_freq, gen.duty_factor,
                                 go.expand_dts([0.02], [12], 12))
            reps += 1
        gait["cpu_numpy_ms"] = (time.perf_counter() - t0) / reps * 1e3
    # ---- the step behind the solve, tau = -J^T f from the result records left on the device
    d_jac = torch.as_tensor(rng.normal(size=(R, 4, 3, 3)), dtype=torch.float64, device=dev)
    d_tau = torch.zeros((R, 4, 3), dtype=torch.float64, device=dev)
    tq_ms = []
    with torch.cuda.stream(stream):
        for it in range(13):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            h.stance_torques(R, d_res.data_ptr(), d_jac.data_ptr(), d_tau.data_ptr())
            e1.record(stream)
            e1.synchronize()
            if it >= 3:
                tq_ms.append(e0.elapsed_time(e1))
    torque = {"workload": f"stance torques [R,4,3] f64 of {R} robots per rank (wb_interface.py:368-371), from the device result records",
              "p50_ms": float(np.median(tq_ms)), "robots_per_s": robots_total / (max_over_ranks(float(np.median(tq_ms))) * 1e-3),
              "algorithmic_bytes": R * (4 * 9 * 8 + 12 * 4 + 12 * 8),
              "note": "3.5 MB of traffic per launch: launch-latency-bound"}
    if rank == 0 and world == 1:
        from oracle import torque_oracle as tq  # CPU leg only
        jac_h, grf_h = d_jac.cpu().numpy(), d_res.cpu().numpy()[:, :12]
        t0 = time.perf_counter()
        tq.stance_torques(jac_h, grf_h)
        torque["cpu_numpy_ms"] = (time.perf_counter() - t0) * 1e3
    h.close()
    return {"workload": f"C5: {robots_total} robots x {n} rollouts, MPPI zero_order (P=144), aliengo trot; {R} robots per rank",
            "gait_clock": gait, "stance_torques": torque,
            "value": robots_total * n / (step_ms * 1e-3), "unit": "rollouts/s", "robot_solves_per_s": robots_total / (step_ms * 1e-3),
            "ms_per_step": step_ms, "scaling": "strong",
            "roofline": {"bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s", "frac": ach / fp32_peak,
                         "traffic": 23960832 // world, "traffic_source": "profiles/r4c_ncu_raw_c5.csv (dram read + write bytes per launch at 8192 robots)",
                         "kernel": "k_rollout<zero_order,fixed,128> (+ k_finalize, one block per robot, in ms_per_step)", "launch_ms": roll_ms, "flop_per_rollout": FLOP_PER_ROLLOUT["zero_order"],
                         "ncu": {"source": "profiles/r4c_ncu_raw_c5.csv (one launch, not this run)", "fma_pipe_pct_of_peak": 39.9, "alu_pipe_pct": 41.3, "xu_pipe_pct": 23.9,
                                 "issue_slots_busy_pct": 71.8, "warp_instructions_per_rollout": 8170, "active_warps_per_scheduler": 2.87}}}


def run_tamols(torch, dev, local_rank, rank, world, barrier, max_over_ranks, with_cpu):
    """TAMOLS foothold scoring (BASELINE.json configs[2] "C3-T" and configs[3] "C4"): footholds/s and candidates/s.
    Device-resident sweep timed with CUDA events on the handle's stream (L2 flushed between launches), the e2e call
    with host buffers (what VisualFootholdAdaptation.compute_adaptation makes) by wall clock, the CPU oracle beside."""
    import synthetic as sy
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    hbm_peak = None
    try:
        hbm_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:  # noqa: BLE001
        pass
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if hbm_peak else "B200_PROFILING.md fallback 6650 GB/s (of fallback)"
    hbm_peak = hbm_peak or 6650.0
    p = tc.make_params(sy.tamols_params_for("go2"), "go2")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {"dtype": "f64", "l2": "flushed (256 MiB write) between timed launches"}

    def timed(h, call, reps):
        stream = torch.cuda.ExternalStream(h.stream, device=dev)
        ms = []
        with torch.cuda.stream(stream):
            for _ in range(3):
                call()
            torch.cuda.synchronize()
            for it in range(reps):
                flush.fill_(it & 0xFF)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                call()
                e1.record(stream)
                e1.synchronize()
                ms.append(e0.elapsed_time(e1))
        return np.array(ms)

    def device_scene(sc):
        d_h = torch.from_numpy(np.ascontiguousarray(sc["heights"])).to(dev)
        d_jobs = torch.from_numpy(np.frombuffer(sc["jobs"].tobytes(), np.uint8).copy()).to(dev)
        d_res = torch.zeros(len(sc["jobs"]) * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        return d_h, d_jobs, d_res

    # ---- C4 sweep: 4 legs x window^2 candidates on one 2048 x 2048 @ 0.02 m map (one robot = the latency regime)
    rng = np.random.default_rng(4 + rank)
    sweep = []
    base = sy.sweep_scene("go2", rng, "stepping_stones_tall", window=256, seed=4)
    d_h = torch.from_numpy(np.ascontiguousarray(base["heights"])).to(dev)
    for w in (32, 64, 128, 256):
        sc = sy.sweep_scene("go2", np.random.default_rng(40 + w), "flat", window=w, map_rows=2048, map_cols=2048, seed=4)
        jobs = sc["jobs"]
        d_jobs = torch.from_numpy(np.frombuffer(jobs.tobytes(), np.uint8).copy()).to(dev)
        d_res = torch.zeros(4 * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        h = tc.Handle(p, device=local_rank, max_jobs=4)
        ms = timed(h, lambda: h.score_device(d_h.data_ptr(), tc.F64, 1, 2048, 2048, 2048, d_jobs.data_ptr(), 4, w, w,
                                             d_res.data_ptr()), 50)
        h.close()
        sweep.append({"window": w, "candidates": 4 * w * w, "p50_ms": float(np.median(ms)),
                      "footholds_per_s": 4 / (np.median(ms) * 1e-3), "candidates_per_s": 4 * w * w / (np.median(ms) * 1e-3)})
    out["sweep_c4"] = sweep

    # ---- batched sweep (the bandwidth-meaningful point): robots x 4 legs x 64 x 64 windows, sharded by robot
    robots_total, w = 2048, 64
    R = robots_total // world
    sc = sy.sweep_scene("go2", np.random.default_rng(90 + rank), "stepping_stones_tall", window=w, seed=4, robots=R)
    d_jobs = torch.from_numpy(np.frombuffer(sc["jobs"].tobytes(), np.uint8).copy()).to(dev)
    d_res = torch.zeros(4 * R * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    h = tc.Handle(p, device=local_rank, max_jobs=4 * R)
    barrier()
    ms = timed(h, lambda: h.score_device(d_h.data_ptr(), tc.F64, 1, 2048, 2048, 2048, d_jobs.data_ptr(), 4 * R, w, w,
                                         d_res.data_ptr()), 20)
    launches = h.launch_count
    # ---- robot states -> seed footholds -> job records -> scores, all on the device (two launches)
    from quadruped_pympc_tamols_b200.helpers.foothold_reference_generator import FootholdReferenceGenerator
    from oracle import seed_oracle as seo  # CPU leg only
    prng = np.random.default_rng(17 + rank)
    pos = np.c_[prng.uniform(-15, 15, (R, 2)), np.full(R, 0.3)]
    eul = np.c_[prng.normal(0, 0.03, (R, 2)), prng.uniform(-3.1, 3.1, R)]
    vel, refv = prng.normal(0, 0.4, (R, 3)), np.c_[prng.normal(0.4, 0.3, R), prng.normal(0, 0.1, R), np.zeros(R)]
    hips = pos[:, None, :] + np.array([[0.19, 0.13, 0.0], [0.19, -0.13, 0.0], [-0.19, 0.13, 0.0], [-0.19, -0.13, 0.0]])
    feet = hips - np.array([0.0, 0.0, 0.3])
    gen = FootholdReferenceGenerator(0.4, np.zeros((R, 4, 3)), 20, 0.3, num_robots=R, handle=h)
    t_in = [torch.as_tensor(a, dtype=torch.float64, device=dev) for a in (pos, eul, vel, refv, hips, np.full(R, 0.3), feet)]
    d_pj = torch.zeros(4 * R * tc.JOB_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    gmap = (2048, 2048, sc["origin"], sc["row_axis"], sc["col_axis"])

    def pipeline():
        gen.compute_footholds_reference(t_in[0], t_in[1], t_in[2], t_in[3], t_in[4], t_in[5], feet_position=t_in[6],
                                        global_map=gmap, window=(w, w), jobs_out=d_pj)
        h.score_device(d_h.data_ptr(), tc.F64, 1, 2048, 2048, 2048, d_pj.data_ptr(), 4 * R, w, w, d_res.data_ptr())

    pms = timed(h, pipeline, 10)
    sms = timed(h, lambda: gen.compute_footholds_reference(t_in[0], t_in[1], t_in[2], t_in[3], t_in[4], t_in[5],
                                                           feet_position=t_in[6], global_map=gmap, window=(w, w),
                                                           jobs_out=d_pj), 10)
    out["pipeline"] = {"workload": f"{R} robots per rank: states -> seed footholds -> tamols_leg_job records -> scores ({w}x{w} windows) on the device",
                       "total_ms": float(np.median(pms)), "seed_jobs_ms": float(np.median(sms)), "launches_per_step": 2}
    if with_cpu:
        t0 = time.perf_counter()
        seo.seed_footholds(pos, eul, vel, refv, hips, np.full(R, 0.3), np.full(R, 0.4), np.zeros((R, 4)),
                           [[] for _ in range(R)], 0.3)
        out["pipeline"]["cpu_seed_oracle_ms"] = (time.perf_counter() - t0) * 1e3
    h.close()
    step_ms = max_over_ranks(float(ms.mean()))
    cand = robots_total * 4 * w * w
    alg_bytes = cand * 8 / world  # each window cell read once per launch on this rank, f64
    ach = alg_bytes / (float(ms.mean()) * 1e-3) / 1e9
    out["batched"] = {"workload": f"{robots_total} robots x 4 legs x {w}x{w} candidate windows on a 2048x2048 f64 map; {R} robots per rank",
                      "ms_per_step": step_ms, "footholds_per_s": robots_total * 4 / (step_ms * 1e-3),
                      "candidates_per_s": cand / (step_ms * 1e-3), "scaling": "strong", "gpu_launches_per_step": 1,
                      "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                                   "traffic": 33991168, "traffic_source": "profiles/r4a_ncu_raw_tamols.csv (dram read + write bytes per launch: the 2048x2048 f64 map once)",
                                   "kernel": "k_tamols_score<double,TMA,32-row tiles>", "launch_ms": float(ms.mean()),
                                   "algorithmic_bytes_per_candidate": 8, "peak_source": peak_src,
                                   "note": "FP64 arithmetic (about 330 FP64 instructions per feasible candidate) binds long before HBM: FP64 pipe 36 % busy, issue slots 62 % (profiles/r4a_ncu_details_tamols.txt); see DESIGN.md"}}
    out["launch_count_check"] = int(launches)

    # ---- e2e: one compute_adaptation-sized call (C3-T: 4 legs x 13 x 7 sensor maps) through tamols_score_host
    rs = np.random.default_rng(300 + rank)
    scenes = [sy.robot_scene("hyqreal1", rs, "stepping_stones") for _ in range(16)]
    ph = tc.make_params(sy.tamols_params_for("hyqreal1"), "hyqreal1")
    h = tc.Handle(ph, device=local_rank, max_jobs=4)
    wall = []
    for it in range(10 + 200):
        s = scenes[it % 16]
        t0 = time.perf_counter()
        h.score_host(s["heights"], s["jobs"], xy=s["xy"])
        if it >= 10:
            wall.append(time.perf_counter() - t0)
    h.close()
    wall = np.array(wall)
    # the same call one level up: VisualFootholdAdaptation.compute_adaptation as wb_interface.py:236-241 makes it
    # (sensor objects in, LegsAttr footholds out; includes the Python packing of the job records)
    from quadruped_pympc_tamols_b200._config import make_config as mk
    from quadruped_pympc_tamols_b200.helpers.legs_attr import LegsAttr
    from quadruped_pympc_tamols_b200.helpers.visual_foothold_adaptation import VisualFootholdAdaptation

    class _Sensor:  # stands in for gym_quadruped's HeightMap: .data[R, C, 1, 3]
        def __init__(self, xy, z):
            self.data = np.concatenate([xy, z[..., None]], axis=-1)[:, :, None, :].copy()

        def get_height(self, p):
            d = self.data[:, :, 0, :].reshape(-1, 3)
            return float(d[np.argmin((d[:, 0] - p[0]) ** 2 + (d[:, 1] - p[1]) ** 2), 2])

    legs = ("FL", "FR", "RL", "RR")
    vfa = VisualFootholdAdaptation(legs, "tamols", config=mk("hyqreal1", tamols_params=sy.tamols_params_for("hyqreal1")),
                                   device=local_rank)
    prepared = []
    for s in scenes:
        prepared.append((LegsAttr(**{n: s["hips"][k] for k, n in enumerate(legs)}),
                         LegsAttr(**{n: _Sensor(s["xy"][k], s["heights"][k]) for k, n in enumerate(legs)}),
                         LegsAttr(**{n: s["feet"][k] for k, n in enumerate(legs)}), s))
    cwall = []
    for it in range(10 + 200):
        hips_, maps_, feet_, s = prepared[it % 16]
        ref_ = LegsAttr(**{n: s["seeds"][k].copy() for k, n in enumerate(legs)})
        vfa.reset()
        t0 = time.perf_counter()
        vfa.compute_adaptation(legs, ref_, hips_, maps_, s["vel"], np.zeros(3), np.zeros(3), base_position=s["base"],
                               current_contact=s["contact"], current_feet_pos=feet_, phase_signal=np.zeros(4))
        if it >= 10:
            cwall.append(time.perf_counter() - t0)
    vfa.close()
    in_b = 4 * 13 * 8 * 8 + 4 * 13 * 7 * 16 + 4 * tc.JOB_DTYPE.itemsize
    out["e2e_c3t"] = {"workload": "C3-T: hyqreal1, 4 legs x 13x7 sensor maps (stepping stones), tamols_score_host",
                      "class_call_p50_ms": float(np.median(cwall) * 1e3),
                      "p50_ms": float(np.median(wall) * 1e3), "footholds_per_s": 4 / float(np.median(wall)),
                      "h2d_bytes_per_step": in_b, "d2h_bytes_per_step": 2 * 4 * tc.RESULT_DTYPE.itemsize,
                      "transfers": "one H2D copy of maps + jobs; results stored by the kernel into mapped pinned host memory as {word, sequence} pairs"}

    if with_cpu:
        from oracle import tamols_oracle as to
        cores = to.max_threads(True)
        t0 = time.perf_counter()
        n = 0
        while time.perf_counter() - t0 < 3.0:
            s = scenes[n % 16]
            to.score(ph, s["heights"], s["jobs"], xy=s["xy"], want_scores=False, fast=True)
            n += 1
        c3t = (time.perf_counter() - t0) / n
        sub = sy.sweep_scene("go2", np.random.default_rng(91), "stepping_stones_tall", window=64, seed=4, robots=32)
        t0, nb = time.perf_counter(), 0
        while time.perf_counter() - t0 < 3.0:
            to.score(p, sub["heights"], sub["jobs"], want_scores=False, fast=True)
            nb += 1
        tb = (time.perf_counter() - t0) / nb
        out["cpu_baseline"] = {"kind": "port", "cores": cores,
                               "c3t_footholds_per_s": 4 / c3t, "c3t_p50_ms": c3t * 1e3,
                               "batched_candidates_per_s": 32 * 4 * 64 * 64 / tb,
                               "sample": f"{n} C3-T calls (3 s) and {nb} batches of 32 robots x 4 legs x 64x64 candidates (3 s), oracle/tamols_oracle.c -O3 OpenMP",
                               "reference_python": "the reference's own Python loop measured 0.75 ms/candidate (15 footholds/s) in the build container (BASELINE.md section 2); it cannot travel to the GPU box"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-batched", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-tamols", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
