"""Tile kernel vs window kernel over the number of jobs (64 x 64 windows on the 2048 x 2048 map): where to switch."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthetic as sy  # noqa: E402
from quadruped_pympc_tamols_b200 import _tamols_cabi as tc  # noqa: E402

dev = torch.device("cuda", 0)
w = int(os.environ.get("W", 64))
p = tc.make_params(sy.tamols_params_for("go2"), "go2")
scene = sy.sweep_scene("go2", np.random.default_rng(90), "stepping_stones_tall", window=w, seed=4, robots=2048)
d_h = torch.from_numpy(np.ascontiguousarray(scene["heights"])).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for robots in (8, 16, 32, 64, 128, 256, 512, 2048):
    jobs = scene["jobs"][: 4 * robots]
    d_jobs = torch.from_numpy(np.frombuffer(jobs.tobytes(), np.uint8).copy()).to(dev)
    d_res = torch.zeros(4 * robots * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    line = f"jobs {4 * robots:5d}:"
    for kern in ("tile", "window"):
        os.environ["TAMOLS_KERNEL"] = kern
        h = tc.Handle(p, max_jobs=4 * robots)
        stream = torch.cuda.ExternalStream(h.stream, device=dev)
        ms = []
        with torch.cuda.stream(stream):
            for it in range(13):
                flush.fill_(it)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                h.score_device(d_h.data_ptr(), tc.F64, 1, 2048, 2048, 2048, d_jobs.data_ptr(), 4 * robots, w, w, d_res.data_ptr())
                e1.record(stream)
                e1.synchronize()
                if it >= 3:
                    ms.append(e0.elapsed_time(e1))
        h.close()
        line += f"  {kern} {np.median(ms) * 1e3:8.1f} us"
    print(line)
