import os, sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
from quadruped_pympc_tamols_b200 import _cabi
inertia = np.diag([0.23, 1.45, 1.5])
for method in ("mppi", "cem_mppi", "random_sampling"):
    outs = []
    for mode in ("", "pair"):
        os.environ["QPMPC_DUMP"] = mode
        h = _cabi.Handle(_cabi.make_config(mass=24.6, inertia=inertia, sampling_method=method, parametrization="zero_order", num_parallel_computations=4097))
        d = torch.empty((4097, h.P), dtype=torch.float32, device="cuda")
        sig = np.linspace(0.5, 4, h.P).astype(np.float32)
        h.generate_noise(42, 0, d.data_ptr(), sigma_in=sig if method == "cem_mppi" else None)
        outs.append(d.cpu().numpy().copy()); h.close()
    a, b = outs
    bad = np.argwhere(a.view(np.uint32) != b.view(np.uint32))
    print(method, "mismatches", len(bad), "of", a.size)
    for i, p in bad[:10]:
        print("  ", i, p, a[i, p], b[i, p], a[i,p].view(np.uint32) - b[i,p].view(np.uint32) if False else (int(a[i,p:p+1].view(np.uint32)[0]) - int(b[i,p:p+1].view(np.uint32)[0])))
