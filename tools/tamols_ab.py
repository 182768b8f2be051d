import sys, os, time
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import synthetic as sy
from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
dev = torch.device("cuda", 0)
R, w = 2048, 64
p = tc.make_params(sy.tamols_params_for("go2"), "go2")
sc = sy.sweep_scene("go2", np.random.default_rng(90), "stepping_stones_tall", window=w, seed=4, robots=R)
d_h = torch.from_numpy(np.ascontiguousarray(sc["heights"])).to(dev)
d_jobs = torch.from_numpy(np.frombuffer(sc["jobs"].tobytes(), np.uint8).copy()).to(dev)
d_res = torch.zeros(4 * R * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
h = tc.Handle(p, max_jobs=4 * R)
stream = torch.cuda.ExternalStream(h.stream, device=dev)
ms = []
with torch.cuda.stream(stream):
    for it in range(13):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        h.score_device(d_h.data_ptr(), tc.F64, 1, 2048, 2048, 2048, d_jobs.data_ptr(), 4 * R, w, w, d_res.data_ptr())
        e1.record(stream); e1.synchronize()
        if it >= 3: ms.append(e0.elapsed_time(e1))
res = np.frombuffer(d_res.cpu().numpy().tobytes(), tc.RESULT_DTYPE)
print(os.environ.get("QPMPC_LIB", "default"), "median ms", np.median(ms), "nfinite mean", res["num_finite"].mean(), "checksum", int(res["best_index"].astype(np.int64).sum()),
      "score bits", hex(int(np.bitwise_xor.reduce(res["best_score"].view(np.uint64)))), "score sum", repr(float(res["best_score"].sum())))
h.close()
