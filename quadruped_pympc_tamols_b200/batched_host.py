"""Batched sampling solve with HOST buffers in and out, copies overlapped with the rollouts (the host
owns the robot states and gathers first GRFs / predicted state / new mean per robot every control tick).

`qpmpc_solve_batched` takes device pointers.  A caller whose robots live in host memory would otherwise do
H2D copy -> launch -> D2H copy one after the other on one stream; at BASELINE.json configs[4] (8192 robots x 1024
rollouts) that is 7.9 MB down and 6.4 MB up around a 1.5 ms launch, a quarter of the step.  Here the robots are cut
into contiguous chunks that alternate between TWO handles (two streams, each with its own scratch): chunk k's inputs
travel while chunk k - 1 integrates and chunk k - 2's results go back -- the copy engines and the SMs all stay busy.
`robot_offset` keeps robot r on noise subsequence r, so the rollouts do not depend on the chunking
(tests/test_batched_host_gpu.py).  Measured at configs[4] on one B200 (tools/c5_e2e_probe.py): 1.74 ms per step with
one chunk, 1.62 / 1.54 / 1.52 / 1.57 ms with 2 / 4 / 8 / 16 -- at 8 chunks the step costs what the launch alone costs.

torch is plumbing here (pinned host memory, device buffers, stream objects around the handles' own streams); the
arithmetic is the library's.  There is no CPU fallback: constructing the solver without a CUDA device raises.
"""
import numpy as np

from . import _cabi
from .sharding import shard_bounds


class BatchedHostSolver:
    """R robots of one configuration.  solve() returns views of pinned host arrays that the next solve() overwrites."""

    def __init__(self, num_robots, chunks=None, device=0, robot_offset=0, **config):
        import torch
        if not torch.cuda.is_available():
            raise _cabi.QpmpcError("BatchedHostSolver needs a CUDA device: this package has no CPU fallback")
        self.torch, self.R, self.robot_offset = torch, int(num_robots), int(robot_offset)
        self.dev = torch.device("cuda", int(device))
        if chunks is None:
            # at most 8 chunks of at least 1,024 robots: a chunk's launch should still fill the GPU a few times over
            # (8 GPUs x 1,024 robots each, 8 chunks of 128: 0.83 ms per step against 0.37 ms unchunked)
            chunks = min(8, self.R // 1024)
        chunks = max(1, min(int(chunks), self.R))
        self.bounds = [shard_bounds(self.R, chunks, k) for k in range(chunks)]
        cap = max(hi - lo for lo, hi in self.bounds)
        n_handles = 2 if chunks > 1 else 1
        self.handles = [_cabi.Handle(_cabi.make_config(max_robots=cap, device=int(device), **config)) for _ in range(n_handles)]
        self.streams = [torch.cuda.ExternalStream(h.stream, device=self.dev) for h in self.handles]
        h0 = self.handles[0]
        self.P, self.H = h0.P, int(config.get("horizon", 12))
        self.cem = config.get("sampling_method") == "cem_mppi"
        self.ga = bool(config.get("gait_adaptive", False))
        R, P, H = self.R, self.P, self.H
        shapes = {"state": (R, 24), "reference": (R, 24), "contact": (R, 4, H), "best": (R, P)}
        if self.cem:
            shapes["sigma"] = (R, P)
        if self.ga:
            shapes.update(timing=(R, 4), nominal=(R,))
        # Host arrays: numpy memory page-locked with cudaHostRegister, NOT torch's caching host allocator -- that one
        # records an event on every stream a block was used on when the block is freed, and the streams here belong to
        # the handles: a view of an output array that outlives close() would abort the process at its last reference.
        # Registered numpy memory stays ordinary memory after cudaHostUnregister, whoever still holds a view.
        self._registered = []
        self.np_in = {k: self._page_locked(s, np.float32) for k, s in shapes.items()}
        self.d_in = {k: torch.zeros(s, dtype=torch.float32, device=self.dev) for k, s in shapes.items()}
        if self.ga:
            self.np_in["oswing"] = self._page_locked((R,), np.int32)
            self.d_in["oswing"] = torch.zeros(R, dtype=torch.int32, device=self.dev)
        out = {"result": (R, _cabi.RESULT_FLOATS), "best": (R, P)}
        if self.cem:
            out["sigma"] = (R, P)
        self.np_out = {k: self._page_locked(s, np.float32) for k, s in out.items()}
        self.d_out = {k: torch.zeros(s, dtype=torch.float32, device=self.dev) for k, s in out.items()}
        self.p_in = {k: torch.from_numpy(v) for k, v in self.np_in.items()}
        self.p_out = {k: torch.from_numpy(v) for k, v in self.np_out.items()}
        if not all(t.is_pinned() for t in list(self.p_in.values()) + list(self.p_out.values())):
            raise _cabi.QpmpcError("cudaHostRegister did not page-lock the host arrays")

    def _page_locked(self, shape, dtype):
        arr = np.zeros(shape, dtype)
        rc = self.torch.cuda.cudart().cudaHostRegister(arr.ctypes.data, arr.nbytes, 0)
        if int(rc) != 0:
            raise _cabi.QpmpcError(f"cudaHostRegister failed with error {int(rc)}")
        self._registered.append(arr)
        return arr

    def io_bytes(self):
        """(host -> device, device -> host) bytes of one solve"""
        return (sum(v.numel() * v.element_size() for v in self.p_in.values()),
                sum(v.numel() * v.element_size() for v in self.p_out.values()))

    def _batch(self, lo, hi):
        d, o = self.d_in, self.d_out
        ptr = lambda t: t[lo:hi].data_ptr()  # noqa: E731  (row slices of contiguous tensors)
        return _cabi.QpmpcBatch(
            num_robots=hi - lo, robot_offset=self.robot_offset + lo, d_state=ptr(d["state"]), d_reference=ptr(d["reference"]),
            d_contact_sequence=ptr(d["contact"]), d_best_in=ptr(d["best"]), d_result=ptr(o["result"]), d_best_out=ptr(o["best"]),
            d_sigma_in=ptr(d["sigma"]) if self.cem else None, d_sigma_out=ptr(o["sigma"]) if self.cem else None,
            d_timing=ptr(d["timing"]) if self.ga else None, d_nominal_freq=ptr(d["nominal"]) if self.ga else None,
            d_optimize_swing=ptr(d["oswing"]) if self.ga else None)

    def solve(self, state, reference, contact_sequence, best, seed, offset, sigma=None, timing=None, nominal_freq=None,
              optimize_swing=None):
        """state [R, 24], reference [R, 24], contact_sequence [R, 4, H], best [R, P] (+ sigma [R, P] for cem_mppi, timing
        [R, 4] / nominal_freq [R] / optimize_swing [R] for the gait-adaptive controllers): numpy arrays or None to keep
        what the pinned input arrays (`self.np_in`) already hold.  Returns (results [R, RESULT_FLOATS], best [R, P]
        [, sigma [R, P]]) as views of the pinned output arrays."""
        torch = self.torch
        given = {"state": state, "reference": reference, "contact": contact_sequence, "best": best, "sigma": sigma,
                 "timing": timing, "nominal": nominal_freq, "oswing": optimize_swing}
        for k, v in given.items():
            if v is not None and k in self.np_in and v is not self.np_in[k]:
                self.np_in[k][...] = v
        for k, (lo, hi) in enumerate(self.bounds):
            h, stream = self.handles[k % len(self.handles)], self.streams[k % len(self.streams)]
            with torch.cuda.stream(stream):
                for name, p in self.p_in.items():
                    self.d_in[name][lo:hi].copy_(p[lo:hi], non_blocking=True)
                h.solve_batched(self._batch(lo, hi), seed, offset)
                for name, p in self.p_out.items():
                    p[lo:hi].copy_(self.d_out[name][lo:hi], non_blocking=True)
        for stream in self.streams:
            stream.synchronize()
        res = (self.np_out["result"], self.np_out["best"])
        return res + ((self.np_out["sigma"],) if self.cem else ())

    def carry_best(self):
        """The mean (and, for cem_mppi, the sigma) the last solve() returned becomes the next solve()'s input without a
        host copy: the page-locked output and input arrays swap roles.  Views returned earlier then alias the INPUTS."""
        for k in ("best", "sigma"):
            if k in self.np_out and k in self.np_in:
                self.np_in[k], self.np_out[k] = self.np_out[k], self.np_in[k]
                self.p_in[k], self.p_out[k] = self.p_out[k], self.p_in[k]

    @property
    def launch_count(self):
        return sum(h.launch_count for h in self.handles)

    def close(self):
        """Idempotent; also runs from __del__ (possibly after a constructor that failed half way)."""
        for stream in getattr(self, "streams", []):
            stream.synchronize()
        for h in getattr(self, "handles", []):
            h.close()
        self.handles, self.streams = [], []
        self.p_in = self.p_out = self.d_in = self.d_out = None
        for arr in getattr(self, "_registered", []):
            self.torch.cuda.cudart().cudaHostUnregister(arr.ctypes.data)
        self._registered = []

    __del__ = close
