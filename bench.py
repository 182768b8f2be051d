#!/usr/bin/env python
"""bench.py -- benchmark of the sampling-MPC hot path and the TAMOLS scoring on B200 (contract: DESIGN.md "Measurement").

Headline workload (BASELINE.json configs[1], "C2"): Go2, bound gait, MPPI, cubic-spline parameters (P = 96),
per-sample step-frequency adaptation, 10,000 rollouts over a 12-step horizon.  One "step" = one MPC solve of that robot.

  value        rollouts/s with the inputs resident in HBM (qpmpc_solve_batched on device tensors), device-timed
               per step with CUDA events on the handle's stream; L2 is flushed between timed steps.
  e2e          the same metric through the reference-facing call qpmpc_solve with HOST buffers, wall clock per call:
               the inputs (1176 B) go host -> device inside the launch (kernel parameter block), the one kernel runs,
               and its finalize step stores the results device -> host into mapped pinned memory as {word, sequence}
               pairs the host waits on (QPMPC_ZEROCOPY=0: H2D copy + kernel + D2H copy + stream sync); p50 reported.
  roofline     the C2 kernel against the FP32 CUDA-core peak measured in this run (no tensor cores: nothing on this
               path is a dense contraction); algorithmic FLOPs per rollout from SURVEY.md 8(d).  roofline.c5 = the
               throughput regime (BASELINE.json configs[4]: 8192 robots x 1024 rollouts, sharded by robot over the ranks,
               no collective), roofline.tamols = the batched TAMOLS scoring.
  configs      every other BASELINE.json config timed the same way: C1 (aliengo, random sampling, zero-order), C3
               (hyqreal, CEM-MPPI, two chained iterations, + one C3-T foothold adaptation = a control tick).
  parity       observed deviations from the CPU oracle on C1, C2, C3 at N = 10,000 (checker only, outside every timed region).
  cpu_baseline the CPU oracle port (-O3, OpenMP over rollouts, all host threads) on the same workload.

`--impl reference` times the reference's CPU implementation of the path.  jax is not installable here, so that arm
is the oracle port of the reference's jax code (kind "port"), all host threads.
"""
import argparse
import csv
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
HYQREAL1 = dict(mass=108.40, inertia=np.array([[4.55031444e+00, 2.75249434e-03, -5.11957307e-01],
                                               [2.75249434e-03, 2.02411774e+01, -7.38560592e-04],
                                               [-5.11957307e-01, -7.38560592e-04, 2.14269772e+01]]))
N_ROLLOUTS = 10000
FLOP_PER_ROLLOUT = {"zero_order": 4380, "linear_spline": 4956, "cubic_spline": 7260, "cubic_spline_ga": 7550}
WORKLOAD = "C2: go2 bound, MPPI, cubic_spline (P=96), per-sample step-frequency adaptation, 10000 rollouts x horizon 12"
PROFILE_TAG = "r6"  # profiles/<tag>_ncu_raw_{c2,c5,tamols}.csv: the ncu captures roofline.traffic is read from

# BASELINE.json configs[0..2]: robot constants, body height, hip offsets, gait (offsets, freq, duty)
CONFIGS = {
    "c1": dict(name="C1: aliengo trot, random_sampling, zero_order (P=144), 10000 rollouts x horizon 12", robot=ALIENGO, h=0.33,
               hip=(0.24, 0.15), gait=([0.5, 1.0, 1.0, 0.5], 1.4, 0.65), method="random_sampling", param="zero_order", ga=False,
               iterations=1, flop="zero_order"),
    "c2": dict(name=WORKLOAD, robot=GO2, h=0.27, hip=(0.19, 0.13), gait=([0.5, 0.5, 0.0, 0.0], 1.8, 0.65), method="mppi",
               param="cubic_spline", ga=True, iterations=1, flop="cubic_spline_ga"),
    "c3": dict(name="C3: hyqreal1 pace, cem_mppi, cubic_spline (P=96), 10000 rollouts x horizon 12, num_sampling_iterations = 2",
               robot=HYQREAL1, h=0.55, hip=(0.44, 0.30), gait=([0.8, 0.3, 0.8, 0.3], 1.4, 0.7), method="cem_mppi",
               param="cubic_spline", ga=False, iterations=2, flop="cubic_spline"),
}


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


def config_problem(key, seed):
    c = CONFIGS[key]
    return synth_problem(np.random.default_rng(seed), c["h"], c["hip"], gait=c["gait"])


def stage(msg):
    """progress on stderr (the JSON line on stdout stays the only stdout output)"""
    print(f"[bench {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def ncu_traffic(what):
    """dram read + write bytes of ONE launch from the committed ncu capture profiles/<tag>_ncu_raw_<what>.csv (row 0
    names, row 1 units, row 2 values), or None when the capture is not there."""
    path = os.path.join(ROOT, "profiles", f"{PROFILE_TAG}_ncu_raw_{what}.csv")
    try:
        rows = list(csv.reader(open(path)))
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        total = 0.0
        for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = rows[0].index(name)
            total += float(rows[2][i]) * scale[rows[1][i]]
        return int(total), os.path.relpath(path, ROOT)
    except Exception:  # noqa: BLE001
        return None, None


def ncu_metrics(what, names):
    """{metric: value} of ONE launch from the committed ncu capture (same file as ncu_traffic), {} when it is absent."""
    path = os.path.join(ROOT, "profiles", f"{PROFILE_TAG}_ncu_raw_{what}.csv")
    try:
        rows = list(csv.reader(open(path)))
        return {k: float(rows[2][rows[0].index(k)]) for k in names if k in rows[0]}
    except Exception:  # noqa: BLE001
        return {}


def c_client_latency(_cabi, key, st, ref, seq, timing, nominal, device, solves=2000, warmup=200):
    """qpmpc_solve timed from a plain C client (tools/c_latency.c, built here with gcc): the C ABI's own end-to-end
    latency with host buffers, without the ctypes call and the Python-side packing that `e2e` includes.  None when
    gcc is missing or anything fails (the figure is an annotation, never the headline)."""
    import shutil
    import subprocess
    import tempfile
    try:
        if shutil.which("gcc") is None:
            return None
        c = CONFIGS[key]
        cfg = _cabi.make_config(sampling_method=c["method"], parametrization=c["param"], gait_adaptive=c["ga"],
                                num_parallel_computations=N_ROLLOUTS, device=device, **c["robot"])
        libdir = os.path.dirname(_cabi.LIB_PATH)
        with tempfile.TemporaryDirectory() as tmp:
            exe, prob = os.path.join(tmp, "c_latency"), os.path.join(tmp, "problem.bin")
            subprocess.run(["gcc", "-std=c99", "-O2", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tools", "c_latency.c"),
                            "-o", exe, "-L", libdir, "-lqpmpc_b200", "-lm", "-Wl,-rpath," + libdir], check=True, capture_output=True)
            P = (c["param"] == "zero_order") * 144 + (c["param"] != "zero_order") * 96
            with open(prob, "wb") as f:
                f.write(bytes(cfg))
                for a in (st, ref, seq, np.zeros(P, np.float32), timing, np.array([nominal, 1.0], np.float32)):
                    f.write(np.ascontiguousarray(a, np.float32).tobytes())
            r = subprocess.run([exe, prob, str(solves), str(warmup)], check=True, capture_output=True, text=True, timeout=120)
        p50, p90, p99, lo = (float(x) for x in r.stdout.split())
        return {"p50_latency_ms": p50 * 1e-3, "p90_latency_ms": p90 * 1e-3, "p99_latency_ms": p99 * 1e-3, "min_latency_ms": lo * 1e-3,
                "solves": solves, "api": "qpmpc_solve called from plain C (tools/c_latency.c), CLOCK_MONOTONIC around each call"}
    except Exception as e:  # noqa: BLE001
        stage(f"C client latency unavailable: {e!r}"[:300])
        return None


# ------------------------------------------------------------------------------------------- reference arm (CPU)
def cpu_solve_loop(key, steps, warmup, seed=0):
    """The CPU port of compute_control_* on one BASELINE config: noise generation + 10k rollouts + update per solve
    (`iterations` of them chained per step, as srbd_controller_interface.py:131-170 does)."""
    from oracle import srbd_oracle as so
    so.set_threads(os.cpu_count() or 1)  # all host threads, also under torchrun (which exports OMP_NUM_THREADS=1)
    c = CONFIGS[key]
    cfg = so.make_cfg(parametrization=c["param"], gait_adaptive=c["ga"], **c["robot"])
    st, ref, seq, timing, nominal = config_problem(key, seed)
    P = so.num_params(cfg)
    best = np.zeros(P, np.float32)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        sigma = np.full(P, 3.0, np.float32)
        for k in range(c["iterations"]):
            eps = so.fill_noise(N_ROLLOUTS, P, 3.0, it * 8 + k, fast=True)
            if c["method"] == "cem_mppi":
                eps = eps * (sigma / 3.0)[None, :]
            freqs = np.asarray([1.4, 2.0, 2.4], np.float32)[(np.arange(N_ROLLOUTS) * 7 + it) % 3] if c["ga"] else None
            r = so.solve(cfg, c["method"], st, ref, seq, best, eps, timing=timing, freqs=freqs, fast=True)
            best = r["best"]
            if c["method"] == "cem_mppi":
                sigma = r["sigma"]
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    return np.array(times), so.max_threads(True)


def line_config():
    """The `config` object: IDENTICAL in both arms (the driver compares them)."""
    return {"workload": WORKLOAD, "l2": "flushed (256 MiB write) between timed steps",
            "multi_gpu": "one independent robot instance per rank, no collective on the control path"}


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    steps = max(args.steps, 200)  # at least 200 timed solves: 20 were warm-up dominated (3.19 M against 3.74 M rollouts/s)
    times, cores = cpu_solve_loop("c2", steps, max(args.warmup, 20))
    ms = float(times.mean() * 1e3)
    value = N_ROLLOUTS / times.mean()
    sample = f"{steps} full solves of the C2 workload (noise + 10000 rollouts + MPPI update each), OpenMP over rollouts"
    line = {"impl": "reference", "metric": "rollouts/sec at 10k samples (one MPPI solve)", "value": value, "unit": "rollouts/s",
            "n_gpus": 1, "ranks_launched": world, "steps": steps, "warmup": max(args.warmup, 20), "ms_per_step": ms,
            "p50_latency_ms": float(np.median(times) * 1e3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": line_config(),
            "notes": "reference jax unavailable on this box; CPU port of the reference math (oracle/srbd_oracle.c, -O3 -fopenmp) on "
                     "rank 0's host cores only: at N > 1 this is ONE host against N GPUs, not a per-GPU ratio",
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


# ------------------------------------------------------------------------------------------- one BASELINE config
class SolveBench:
    """One of C1 / C2 / C3 on one GPU: the device-resident solve (CUDA events) and the host-buffer call (wall clock)."""

    def __init__(self, torch, _cabi, dev, local_rank, key, seed):
        self.torch, self._cabi, self.dev, self.key = torch, _cabi, dev, key
        c = self.c = CONFIGS[key]
        self.st, self.ref, self.seq, self.timing, self.nominal = config_problem(key, seed)
        self.h = _cabi.Handle(_cabi.make_config(sampling_method=c["method"], parametrization=c["param"], gait_adaptive=c["ga"],
                                                num_parallel_computations=N_ROLLOUTS, device=local_rank, **c["robot"]))
        P = self.P = self.h.P
        self.stream = torch.cuda.ExternalStream(self.h.stream, device=dev)
        t_ = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=dev)  # noqa: E731
        self.d = dict(state=t_(self.st[None]), ref=t_(self.ref[None]), seq=t_(self.seq[None]), timing=t_(self.timing[None]),
                      nom=t_(np.array([self.nominal])), osw=t_(np.array([1]), torch.int32),
                      best=[torch.zeros((1, P), device=dev), torch.zeros((1, P), device=dev)],
                      sigma=[torch.full((1, P), 3.0, device=dev), torch.full((1, P), 3.0, device=dev)],
                      res=torch.zeros((1, _cabi.RESULT_FLOATS), device=dev))
        self.cem = c["method"] == "cem_mppi"

    def device_step(self, it):
        """One control tick with everything resident: `iterations` launches back to back (ping-pong mean / sigma)."""
        d, _cabi = self.d, self._cabi
        for k in range(self.c["iterations"]):
            j = it * self.c["iterations"] + k
            if self.cem and k == 0:
                d["sigma"][j & 1].fill_(3.0)  # srbd_controller_interface.py:129: sigma restarts at 3 every tick
            b = _cabi.QpmpcBatch(num_robots=1, d_state=d["state"].data_ptr(), d_reference=d["ref"].data_ptr(),
                                 d_contact_sequence=d["seq"].data_ptr(), d_best_in=d["best"][j & 1].data_ptr(),
                                 d_timing=d["timing"].data_ptr(), d_nominal_freq=d["nom"].data_ptr(),
                                 d_optimize_swing=d["osw"].data_ptr(), d_result=d["res"].data_ptr(),
                                 d_best_out=d["best"][(j + 1) & 1].data_ptr(),
                                 d_sigma_in=d["sigma"][j & 1].data_ptr() if self.cem else None,
                                 d_sigma_out=d["sigma"][(j + 1) & 1].data_ptr() if self.cem else None)
            self.h.solve_batched(b, 42, j)

    def time_device(self, K, W, flush, clocks=None, barrier=None):
        torch = self.torch
        with torch.cuda.stream(self.stream):
            for it in range(W):
                self.device_step(it)
            (barrier or torch.cuda.synchronize)()
            launches0 = self.h.launch_count
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
            if clocks:
                clocks.sample()
            for it in range(K):
                flush.fill_(it & 0xFF)  # untimed L2 flush between timed steps
                ev[it][0].record(self.stream)
                self.device_step(W + it)
                ev[it][1].record(self.stream)
                if clocks and it % max(1, K // 8) == 0:
                    clocks.sample()
            if clocks:
                clocks.sample()
            (barrier or torch.cuda.synchronize)()
        return np.array([a.elapsed_time(b) for a, b in ev]), self.h.launch_count - launches0

    def event_floor(self, K, flush):
        """What the device-timing method itself reads for a kernel that does nothing: the same event pair on the same
        stream, the same L2 flush ahead of it, around a one-block torch kernel (launch + event record latencies)."""
        torch = self.torch
        one = torch.zeros(32, device=self.dev)
        with torch.cuda.stream(self.stream):
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
            for it in range(K + 10):
                flush.fill_(it & 0xFF)
                if it >= 10:
                    ev[it - 10][0].record(self.stream)
                one.add_(1.0)
                if it >= 10:
                    ev[it - 10][1].record(self.stream)
            torch.cuda.synchronize()
        return float(np.median([a.elapsed_time(b) for a, b in ev]))

    def time_e2e(self, K, W):
        """qpmpc_solve / qpmpc_solve_iterated with host buffers, wall clock around the Python-level call."""
        best = np.zeros(self.P, np.float32)
        sig = np.full(self.P, 3.0, np.float32) if self.cem else None
        ts = []
        for it in range(W + K):
            t0 = time.perf_counter()
            r = self.h.solve(self.st, self.ref, self.seq, best, 42, it * self.c["iterations"], timing=self.timing,
                             nominal_freq=self.nominal, optimize_swing=1, sigma_in=sig, want_costs=False,
                             iterations=self.c["iterations"])
            dt_ = time.perf_counter() - t0
            best = r["best"]
            if it >= W:
                ts.append(dt_)
        return np.array(ts)

    def io_bytes(self):
        P = self.P
        return 4 * (24 + 24 + 4 * 12 + 2 * P + 4 + 2), 8 * (self._cabi.RESULT_FLOATS + P * (2 if self.cem else 1))

    def close(self):
        self.h.close()


def parity_block(torch, _cabi, dev, local_rank):
    """Observed deviations of the CUDA path from the CPU oracle on C1, C2, C3 at N = 10,000 with the SAME noise (dumped
    by qpmpc_generate_noise).  The oracle is the checker here, never the thing timed.  Bars (BASELINE.json): costs 1e-4
    relative, updated control sequence 1e-3; tests/util.py:softmin_tolerance widens the latter by the softmin's
    conditioning -- `plain_1e-3_held` says whether that was needed at all."""
    from oracle import srbd_oracle as so
    from tests.util import rel_err, softmin_tolerance, softmin_update_f64
    out = {}
    for key, c in CONFIGS.items():
        st, ref, seq, timing, nominal = config_problem(key, 4242)
        kw = dict(parametrization=c["param"], gait_adaptive=c["ga"], **c["robot"])
        h = _cabi.Handle(_cabi.make_config(sampling_method=c["method"], num_parallel_computations=N_ROLLOUTS, device=local_rank, **kw))
        ocfg = so.make_cfg(**kw)
        P = h.P
        best = (2.0 * np.random.default_rng(7).normal(size=P)).astype(np.float32)
        sigma = np.full(P, 3.0, np.float32)
        d_eps = torch.empty((N_ROLLOUTS, P), dtype=torch.float32, device=dev)
        d_freqs = torch.empty(N_ROLLOUTS, dtype=torch.float32, device=dev)
        rows = []
        for it in range(c["iterations"]):
            cem = c["method"] == "cem_mppi"
            h.generate_noise(42, it, d_eps.data_ptr(), d_freqs.data_ptr(), sigma_in=sigma if cem else None, nominal_freq=nominal,
                             optimize_swing=1)
            eps = d_eps.cpu().numpy()
            freqs = d_freqs.cpu().numpy() if c["ga"] else None
            r = h.solve(st, ref, seq, best, 42, it, timing=timing, nominal_freq=nominal, optimize_swing=1, sigma_in=sigma if cem else None)
            o = so.solve(ocfg, c["method"], st, ref, seq, best, eps, timing=timing, freqs=freqs, fast=False)
            co = np.asarray(o["costs"], np.float64)
            w = np.exp(-(co - co.min()))
            w /= w.sum()
            live = w > 1e-7
            row = {"iteration": it, "kernel": h.last_kernel, "cost_max_rel_err": rel_err(r["costs"], o["costs"]),
                   "best_max_abs_err": float(np.max(np.abs(r["best"] - o["best"]))),
                   "grf_max_abs_err": float(np.max(np.abs(r["grf"] - o["grf"]))),
                   "best_index_equal": bool(r["best_index"] == o["best_index"])}
            if c["method"] == "random_sampling":
                row.update(tolerance_applied=1e-3, delta=None, live_samples=None)
            else:
                row.update(tolerance_applied=float(softmin_tolerance(r["costs"], o["costs"], eps)),
                           delta=float(np.max(np.abs(np.asarray(r["costs"], np.float64)[live] - co[live]))),
                           live_samples=int(live.sum()),
                           update_vs_f64_softmin_of_own_costs=float(np.max(np.abs(r["best"] - softmin_update_f64(best, r["costs"], eps)))))
            row["plain_1e-3_held"] = bool(row["best_max_abs_err"] < 1e-3)
            row["cost_bar_1e-4_held"] = bool(row["cost_max_rel_err"] < 1e-4)
            if cem:
                row["sigma_max_rel_err"] = rel_err(r["sigma"], o["sigma"])
                sigma = r["sigma"]
            rows.append(row)
            best = r["best"]
        h.close()
        out[key] = {"workload": c["name"], "iterations": rows}
    return out


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

    K, W = max(args.steps, 1), max(args.warmup, 3)
    with_cpu = rank == 0 and world == 1 and not args.no_cpu
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    clocks = Clocks(local_rank)

    # ---- headline: C2, one independent robot instance per rank (batched_simulations.py style)
    stage("C2 device-resident")
    c2 = SolveBench(torch, _cabi, dev, local_rank, "c2", 1000 + rank)
    h, P = c2.h, c2.P
    dev_ms, launches = c2.time_device(K, W, flush, clocks, barrier)
    total_ms = max_over_ranks(float(dev_ms.sum()))
    value = N_ROLLOUTS * K * world / (total_ms * 1e-3)
    c2_kernel = h.last_kernel
    floor_ms = c2.event_floor(min(K, 100), flush)

    # roofline of the dominant kernel: events around each kernel inside the solve
    h.set_profiling(True)
    roll = []
    with torch.cuda.stream(c2.stream):
        for it in range(min(K, 200)):
            flush.fill_(it & 0xFF)
            c2.device_step(W + K + it)
            roll.append(h.profile_last())
    h.set_profiling(False)
    roll = np.array(roll)
    fp32_peak = _cabi.measure_fp32_peak(local_rank)
    ach = FLOP_PER_ROLLOUT["cubic_spline_ga"] * N_ROLLOUTS / (roll[:, 0].mean() * 1e-3) / 1e12

    stage("C2 e2e")
    e2e = c2.time_e2e(K, W)
    e2e_total = max_over_ranks(float(e2e.sum()))
    e2e_value = N_ROLLOUTS * K * world / e2e_total
    in_bytes, out_bytes = c2.io_bytes()
    st, ref, seq, timing, nominal = c2.st, c2.ref, c2.seq, c2.timing, c2.nominal
    best = np.zeros(P, np.float32)
    c_client = c_client_latency(_cabi, "c2", st, ref, seq, timing, nominal, local_rank) if rank == 0 else None

    # ---- num_sampling_iterations = 3 chained on the device (qpmpc_solve_iterated) against three
    # host-level calls (the upstream interface loop)
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

    stage("C2 class call / iterated")
    class_p50 = timed(class_call)
    ctl.close()
    iterated = {"workload": "C2 with num_sampling_iterations = 3 (srbd_controller_interface.py:131-170)",
                "chained_p50_ms": timed(lambda it: h.solve(st, ref, seq, best, 42, 3 * it, timing=timing, nominal_freq=nominal,
                                                           optimize_swing=1, want_costs=False, iterations=3)),
                "three_calls_p50_ms": timed(three_calls)}

    # ---- the other BASELINE configs, timed the same way (rank-local; the CPU port beside them on rank 0 at N = 1)
    configs = {}
    tamols = None
    if not args.no_tamols:
        stage("TAMOLS")
        tamols = run_tamols(torch, dev, local_rank, rank, world, barrier, max_over_ranks, with_cpu=with_cpu)
    for key in ("c1", "c3"):
        stage(key)
        sb = SolveBench(torch, _cabi, dev, local_rank, key, 2000 + rank)
        kd, kl = sb.time_device(min(K, 100), W, flush)
        ke = sb.time_e2e(min(K, 100), W)
        ib, ob = sb.io_bytes()
        its = CONFIGS[key]["iterations"]
        entry = {"workload": CONFIGS[key]["name"], "kernel": sb.h.last_kernel, "launches_per_step": int(kl // min(K, 100)),
                 "device_p50_ms": float(np.median(kd)), "e2e_p50_ms": float(np.median(ke)), "e2e_p99_ms": float(np.percentile(ke, 99)),
                 "rollouts_per_s_device": N_ROLLOUTS * its / (float(np.median(kd)) * 1e-3),
                 "rollouts_per_s_e2e": N_ROLLOUTS * its / float(np.median(ke)),
                 "h2d_bytes_per_step": ib, "d2h_bytes_per_step": ob,
                 "api": "qpmpc_solve" if its == 1 else f"qpmpc_solve_iterated (iterations = {its}, one wait)"}
        if key == "c3" and tamols is not None:
            # one control tick of configs[2]: the CEM solve followed by one foothold adaptation (C3-T, 4 legs x 13 x 7)
            entry["control_tick_e2e_p50_ms"] = float(np.median(ke) * 1e3 + tamols["e2e_c3t"]["p50_ms"])
            entry["control_tick"] = "qpmpc_solve_iterated(2) + tamols_score_host (C3-T), p50 of each, summed"
        entry["e2e_p50_ms"] *= 1e3
        entry["e2e_p99_ms"] *= 1e3
        if with_cpu:
            ct, cores = cpu_solve_loop(key, 300, 5)
            entry["cpu_baseline"] = {"kind": "port", "cores": cores, "p50_ms": float(np.median(ct) * 1e3),
                                     "rollouts_per_s": N_ROLLOUTS * its / float(ct.mean()),
                                     "sample": f"300 control ticks ({its} solve(s) each: noise + 10000 rollouts + update), {ct.sum():.1f} s of CPU work"}
        sb.close()
        configs[key] = entry

    # ---- batched robots C5 (sharded by robot, no collective)
    batched = None
    if not args.no_batched:
        stage("C5 batched")
        batched = run_batched(torch, _cabi, dev, local_rank, rank, world, barrier, max_over_ranks, fp32_peak, with_cpu)

    parity = None
    if rank == 0 and world == 1 and not args.no_parity:
        stage("parity block")
        parity = parity_block(torch, _cabi, dev, local_rank)

    # ---- CPU baseline (rank 0, N=1 only): bounded sample of the same workload
    cpu = None
    if with_cpu:
        stage("C2 CPU baseline")
        nsolve = 4000  # about 10 s of CPU work on the B200 host (2.4 ms per solve with 16 threads)
        times, cores = cpu_solve_loop("c2", nsolve, 20)
        cpu = {"value": N_ROLLOUTS / float(times.mean()), "unit": "rollouts/s", "cores": cores, "kind": "port",
               "sample": f"{nsolve} full C2 solves (noise + 10000 rollouts + MPPI update), {times.sum():.1f} s of CPU work",
               "p50_latency_ms": float(np.median(times) * 1e3)}

    clk = clocks.summary()
    if rank == 0:
        traffic, tsrc = ncu_traffic("c2")
        roofline = {"bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s", "frac": ach / fp32_peak,
                    "traffic": traffic, "traffic_source": tsrc and f"dram__bytes_read.sum + dram__bytes_write.sum of one launch, {tsrc}",
                    "kernel": f"k_rollout_{c2_kernel}<cubic,gait_adaptive> (rollouts + fused finalize, one launch per solve)",
                    "launch_ms": float(roll[:, 0].mean()), "finalize_ms": float(roll[:, 1].mean()),
                    "flop_per_rollout": FLOP_PER_ROLLOUT["cubic_spline_ga"],
                    "peak_source": "FP32 FMA-chain microbenchmark in this run (qpmpc_measure_fp32_peak, constant-operand FFMA); MEASURED_PEAKS.json has no FP32 figure",
                    "note": "10k rollouts = 75.5 MFLOP: latency-bound at this size; roofline.c5 is the throughput regime",
                    "c5": batched and batched["roofline_c5"], "tamols": tamols and tamols["batched"]["roofline"]}
        line = {"metric": "rollouts/sec at 10k samples (one MPPI solve)", "value": value, "unit": "rollouts/s", "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": float(total_ms / K), "p50_latency_ms": float(np.median(dev_ms)),
                "device_timing": {"event_pair_floor_p50_ms": floor_ms,
                                  "solve_p50_minus_floor_ms": float(np.median(dev_ms)) - floor_ms,
                                  "note": "floor = the same CUDA-event pair, stream and L2 flush around a one-block kernel: launch + "
                                          "event-record latency that every device-timed step carries; value / ms_per_step include it"},
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": line_config(),
                "notes": {"noise": "in-kernel: one Philox4x32-7 seed per 24 values + xoroshiro64+ bursts of 12, Box-Muller on the SFU (stream v3)",
                          "scaling": "the headline is one C2 robot per rank (weak); roofline.c5 carries the sharded 8192-robot step (strong) at this N"},
                "e2e": {"value": e2e_value, "unit": "rollouts/s", "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": out_bytes,
                        "p50_latency_ms": float(np.median(e2e) * 1e3), "p99_latency_ms": float(np.percentile(e2e, 99) * 1e3),
                        "class_call_p50_latency_ms": class_p50, "c_client": c_client,
                        "api": "qpmpc_solve (host buffers; costs[N] left on the device as upstream)",
                        "transfers": "inputs in the kernel parameter block (host -> device with the launch); results stored by the kernel into mapped pinned host memory as {word, sequence} pairs",
                        "c5": batched and batched["e2e"]},
                "gpu_launches": int(launches),
                "roofline": roofline, "cpu_baseline": cpu, "clocks": clk, "configs": configs, "parity": parity,
                "iterated": iterated, "batched": batched, "tamols": tamols}
        print(json.dumps(line))
    c2.close()
    if world > 1:
        dist.destroy_process_group()


def run_batched(torch, _cabi, dev, local_rank, rank, world, barrier, max_over_ranks, fp32_peak, with_cpu,
                robots_total=8192, n=1024):
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
    h_state = (np.stack([probs[i][0] for i in idx]) + rng.normal(0, 1e-3, (R, 24))).astype(np.float32)
    h_ref, h_seq = np.stack([probs[i][1] for i in idx]), np.stack([probs[i][2] for i in idx])
    d_state, d_ref, d_seq = t_(h_state), t_(h_ref), t_(h_seq)
    d_best = [torch.zeros((R, P), device=dev), torch.zeros((R, P), device=dev)]
    d_res = torch.zeros((R, _cabi.RESULT_FLOATS), device=dev)
    stream = torch.cuda.ExternalStream(h.stream, device=dev)
    h.set_profiling(True)
    steps, roll = 10, []

    def batch(it):
        return _cabi.QpmpcBatch(num_robots=R, robot_offset=lo, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(),
                                d_contact_sequence=d_seq.data_ptr(), d_best_in=d_best[it & 1].data_ptr(),
                                d_result=d_res.data_ptr(), d_best_out=d_best[(it + 1) & 1].data_ptr())

    with torch.cuda.stream(stream):
        for it in range(3 + steps):
            if it == 3:
                barrier()
            h.solve_batched(batch(it), 7, it)
            if it >= 3:
                roll.append(h.profile_last())
        barrier()
    roll = np.array(roll)
    kernel = h.last_kernel
    step_ms = max_over_ranks(float((roll[:, 0] + roll[:, 1]).mean()))
    roll_ms = max_over_ranks(float(roll[:, 0].mean()))
    ach = FLOP_PER_ROLLOUT["zero_order"] * R * n / (roll_ms * 1e-3) / 1e12
    h.set_profiling(False)

    stage("C5 e2e")
    # ---- C5 end to end (the host owns the robot states and gathers GRFs / predicted state / mean per
    # robot): pinned host arrays -> H2D of state, reference, contact sequence and previous mean -> one launch -> D2H of
    # the result records and the new mean; wall clock per step including the final stream synchronise
    p_in = [torch.from_numpy(a).pin_memory() for a in (h_state, h_ref.astype(np.float32), h_seq.astype(np.float32),
                                                       np.zeros((R, P), np.float32))]
    p_res = torch.zeros((R, _cabi.RESULT_FLOATS)).pin_memory()
    p_best = torch.zeros((R, P)).pin_memory()
    wall = []
    with torch.cuda.stream(stream):
        for it in range(3 + steps):
            if it == 3:
                barrier()
            t0 = time.perf_counter()
            d_state.copy_(p_in[0], non_blocking=True)
            d_ref.copy_(p_in[1], non_blocking=True)
            d_seq.copy_(p_in[2], non_blocking=True)
            d_best[it & 1].copy_(p_in[3], non_blocking=True)
            h.solve_batched(batch(it), 7, 100 + it)
            p_res.copy_(d_res, non_blocking=True)
            p_best.copy_(d_best[(it + 1) & 1], non_blocking=True)
            stream.synchronize()
            p_in[3].copy_(p_best)  # the host carries the mean to the next tick
            if it >= 3:
                wall.append(time.perf_counter() - t0)
        barrier()
    # the pinned blocks must go back to torch's host allocator BEFORE the handle (and with it the stream they were used
    # on) is destroyed: the allocator records an event on that stream when a block is freed
    del p_in, p_res, p_best
    serial_s = max_over_ranks(float(np.mean(wall)))
    # the same step through the package's host-buffer solver: robots in up to 8 chunks (at least 1,024 robots each) alternating between two handles /
    # streams, chunk k's copies overlapping chunk k - 1's rollouts (quadruped_pympc_tamols_b200/batched_host.py)
    from quadruped_pympc_tamols_b200.batched_host import BatchedHostSolver
    hs = BatchedHostSolver(R, device=local_rank, robot_offset=lo, sampling_method="mppi", parametrization="zero_order",
                           num_parallel_computations=n, **ALIENGO)
    hs.np_in["state"][...], hs.np_in["reference"][...], hs.np_in["contact"][...] = h_state, h_ref, h_seq
    wall = []
    for it in range(3 + steps):
        if it == 3:
            barrier()
        t0 = time.perf_counter()
        hs.solve(None, None, None, None, 7, 100 + it)
        hs.carry_best()  # the host carries the mean to the next tick: the page-locked output array is the next input
        if it >= 3:
            wall.append(time.perf_counter() - t0)
    barrier()
    h2d_b, d2h_b = hs.io_bytes()
    hs_chunks = len(hs.bounds)
    hs_launches = hs.launch_count // (3 + steps)
    hs.close()
    e2e_s = max_over_ranks(float(np.mean(wall)))
    e2e = {"workload": f"C5 through host buffers: {robots_total} robots x {n} rollouts; {R} robots per rank",
           "value": robots_total * n / e2e_s, "unit": "rollouts/s", "ms_per_step": e2e_s * 1e3,
           "h2d_bytes_per_step": int(h2d_b), "d2h_bytes_per_step": int(d2h_b), "gpu_launches_per_step": int(hs_launches),
           "api": f"BatchedHostSolver.solve: page-locked host arrays, {hs_chunks} chunk(s) of robots alternating between two handles / streams "
                  "(H2D copy, qpmpc_solve_batched, D2H copy per chunk), one synchronise per stream and step",
           "serial_ms_per_step": serial_s * 1e3,
           "serial_api": "one stream: H2D copies -> one qpmpc_solve_batched launch over all robots -> D2H copies -> synchronise"}

    stage("C5 cpu / gait / torque")
    cpu = None
    if with_cpu:  # the CPU port on a 64-robot subset of the same workload, all host threads over the rollouts of one robot at a time
        from oracle import srbd_oracle as so
        so.set_threads(os.cpu_count() or 1)
        ocfg = so.make_cfg(parametrization="zero_order", **ALIENGO)
        sub, t0 = 64, time.perf_counter()
        for r_ in range(sub):
            eps = so.fill_noise(n, P, 3.0, r_, fast=True)
            so.solve(ocfg, "mppi", h_state[r_], h_ref[r_], h_seq[r_], np.zeros(P, np.float32), eps, fast=True)
        t_sub = time.perf_counter() - t0
        cpu = {"kind": "port", "cores": so.max_threads(True), "value": sub * n / t_sub, "unit": "rollouts/s",
               "ms_per_step_extrapolated": t_sub / sub * robots_total * 1e3,
               "sample": f"{sub} of the {robots_total} robots (noise + {n} rollouts + MPPI update each), {t_sub:.2f} s; the full-step time is EXTRAPOLATED x{robots_total // sub}"}

    # ---- the step in front of the rollouts, contact sequences of this rank's robots from R gait clocks
    from quadruped_pympc_tamols_b200.helpers.periodic_gait_generator import PeriodicGaitGenerator
    gen = PeriodicGaitGenerator(rng.uniform(0.5, 0.8, R), rng.uniform(1.0, 2.4, R), rng.integers(0, 7, R), 12,
                                num_robots=R, handle=h)
    ph0 = rng.uniform(0, 1, (R, 4))
    gen.set_phase_signal(ph0)

    def event_timed(call, reps=10):
        ms = []
        with torch.cuda.stream(stream):
            for it in range(3 + reps):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                call()
                e1.record(stream)
                e1.synchronize()
                if it >= 3:
                    ms.append(e0.elapsed_time(e1))
        return ms

    gait_ms = event_timed(lambda: gen.compute_contact_sequence([0.02], [12]))
    gait = {"workload": f"contact sequences [R,4,12] of {R} gait clocks per rank (PeriodicGaitGenerator.compute_contact_sequence)",
            "p50_ms": float(np.median(gait_ms)), "sequences_per_s": robots_total / (max_over_ranks(float(np.median(gait_ms))) * 1e-3),
            "algorithmic_bytes": R * (4 * 8 + 4 + 4 * 8 + 8 + 8 + 1 + 4 * 12 * 4),
            "note": "2.3 MB of traffic per launch: launch-latency-bound, HBM time at the measured peak is 0.35 us"}
    if with_cpu:
        from oracle import gait_oracle as go  # CPU leg only
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 2.0:
            go.contact_sequences(ph0, np.zeros((R, 4), bool), gen.phase_offset, gen.step_freq, gen.duty_factor,
                                 go.expand_dts([0.02], [12], 12))
            reps += 1
        gait["cpu_numpy_ms"] = (time.perf_counter() - t0) / reps * 1e3
    # ---- the steps behind the solve: tau = -J^T f and the observables / dataset rows, from the result
    # records left on the device
    d_jac = torch.as_tensor(rng.normal(size=(R, 4, 3, 3)), dtype=torch.float64, device=dev)
    d_tau = torch.zeros((R, 4, 3), dtype=torch.float64, device=dev)
    tq_ms = event_timed(lambda: h.stance_torques(R, d_res.data_ptr(), d_jac.data_ptr(), d_tau.data_ptr()))
    torque = {"workload": f"stance torques [R,4,3] f64 of {R} robots per rank (wb_interface.py:368-371), from the device result records",
              "p50_ms": float(np.median(tq_ms)), "robots_per_s": robots_total / (max_over_ranks(float(np.median(tq_ms))) * 1e-3),
              "algorithmic_bytes": R * (4 * 9 * 8 + 12 * 4 + 12 * 8),
              "note": "3.5 MB of traffic per launch: launch-latency-bound"}
    if with_cpu:
        from oracle import torque_oracle as tq  # CPU leg only
        jac_h, grf_h = d_jac.cpu().numpy(), d_res.cpu().numpy()[:, :12]
        t0 = time.perf_counter()
        tq.stance_torques(jac_h, grf_h)
        torque["cpu_numpy_ms"] = (time.perf_counter() - t0) * 1e3
    h.close()
    traffic, tsrc = ncu_traffic("c5")
    kname = {"pair": "k_rollout_pair<mppi> (two rollouts per thread, packed FP32, one block per robot, update fused)",
             "thread": "k_rollout<zero_order,fixed,128> (+ k_finalize)"}.get(kernel, kernel)
    roofline_c5 = {"bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s", "frac": ach / fp32_peak,
                   "frac_whole_step": FLOP_PER_ROLLOUT["zero_order"] * R * n / (step_ms * 1e-3) / 1e12 / fp32_peak,
                   "traffic": traffic and traffic // world, "traffic_source": tsrc and f"{tsrc} (dram read + write bytes per launch at {robots_total} robots)",
                   "kernel": kname, "launch_ms": roll_ms, "ms_per_step": step_ms, "n_gpus": world, "robots_per_rank": R,
                   "rollouts_per_s": robots_total * n / (step_ms * 1e-3), "scaling": "strong",
                   "flop_per_rollout": FLOP_PER_ROLLOUT["zero_order"],
                   "note": "the timed steps continue one MPPI iteration chain on a fixed state: the cost spread shrinks step by step and "
                           "the number of live samples whose rows are drawn again grows (launch_ms is the mean over steps 3..12)"}
    return {"workload": f"C5: {robots_total} robots x {n} rollouts, MPPI zero_order (P=144), aliengo trot; {R} robots per rank",
            "gait_clock": gait, "stance_torques": torque,
            "value": robots_total * n / (step_ms * 1e-3), "unit": "rollouts/s", "robot_solves_per_s": robots_total / (step_ms * 1e-3),
            "ms_per_step": step_ms, "scaling": "strong", "e2e": e2e, "cpu_baseline": cpu, "roofline_c5": roofline_c5}


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
    tamols_kernel = {"window": "k_tamols_window<double,128> (one block per job, whole window)",
                     "tile32": "k_tamols_score<double,TMA,32-row tiles>", "tile8": "k_tamols_score<double,TMA,8-row tiles>"}[h.last_kernel]
    # ---- robot states -> seed footholds -> job records -> scores, all on the device (two launches)
    from quadruped_pympc_tamols_b200.helpers.foothold_reference_generator import FootholdReferenceGenerator
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
        from oracle import seed_oracle as seo  # CPU leg only
        t0 = time.perf_counter()
        seo.seed_footholds(pos, eul, vel, refv, hips, np.full(R, 0.3), np.full(R, 0.4), np.zeros((R, 4)),
                           [[] for _ in range(R)], 0.3)
        out["pipeline"]["cpu_seed_oracle_ms"] = (time.perf_counter() - t0) * 1e3
    h.close()
    step_ms = max_over_ranks(float(ms.mean()))
    cand = robots_total * 4 * w * w
    alg_bytes = cand * 8 / world  # each window cell read once per launch on this rank, f64
    ach = alg_bytes / (float(ms.mean()) * 1e-3) / 1e9
    traffic, tsrc = ncu_traffic("tamols")
    out["batched"] = {"workload": f"{robots_total} robots x 4 legs x {w}x{w} candidate windows on a 2048x2048 f64 map; {R} robots per rank",
                      "ms_per_step": step_ms, "footholds_per_s": robots_total * 4 / (step_ms * 1e-3),
                      "candidates_per_s": cand / (step_ms * 1e-3), "scaling": "strong", "gpu_launches_per_step": 1,
                      "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                                   "traffic": traffic, "traffic_source": tsrc and f"{tsrc} (dram read + write bytes per launch: the 2048x2048 f64 map once)",
                                   "kernel": tamols_kernel, "launch_ms": float(ms.mean()),
                                   "ncu": dict(ncu_metrics("tamols", ["sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
                                                                      "smsp__issue_active.avg.pct_of_peak_sustained_active",
                                                                      "smsp__inst_executed.sum"]),
                                               source=f"profiles/{PROFILE_TAG}_ncu_raw_tamols.csv (one launch under ncu --set full)"),
                                   "ms_per_step": step_ms, "n_gpus": world, "candidates_per_s": cand / (step_ms * 1e-3),
                                   "algorithmic_bytes_per_candidate": 8, "peak_source": peak_src,
                                   "note": "FP64 arithmetic binds long before HBM (windows overlap in L2: measured dram traffic is 8x below the algorithmic bytes): the binding resource is in `ncu` (FP64 pipe and issue slots, % of peak); see DESIGN.md"}}
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
                               "note": "at the reference's own size (4 x 91 candidates) one GPU call and the OpenMP C port both take about 30 us: "
                                       "the call is a launch-latency floor, the GPU case is the batched one",
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
    ap.add_argument("--no-parity", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
