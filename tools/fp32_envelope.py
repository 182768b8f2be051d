#!/usr/bin/env python
"""How far apart do faithful float32 implementations of the reference's sampling path sit?  

The fixtures pin this repo's path to the reference's source executed under a numpy stand-in; XLA's own code generation
(FMA contraction, its transcendentals) cannot be run here.  This tool quantifies that gap from the other side: the C
restatement is compiled three ways --
    strict     -O2 -ffp-contract=off, libm sinf / cosf                      (the checker)
    contract   -O3 -march=x86-64-v3 -ffp-contract=fast (FMAs wherever the compiler finds them), libm sinf / cosf
    alt-trig   strict arithmetic, sin / cos evaluated in double and rounded to float
-- and every sampling fixture is solved by each; the pairwise spread of the costs (max relative) and of the updated
control sequence (max absolute) is what ANY two faithful float32 implementations exhibit.  With --gpu the CUDA path's
deviation from the strict build is printed next to it (run on the GPU box).
Usage: python tools/fp32_envelope.py [--gpu] > profiles/r5_fp32_envelope.txt"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import srbd_oracle as so  # noqa: E402
from tests.util import fixture_kwargs, rel_err, sampling_fixtures  # noqa: E402

BUILD = os.path.join(ROOT, "oracle", "_build")
VARIANTS = {"contract": ["-O3", "-march=x86-64-v3", "-ffp-contract=fast"],
            "alt-trig": ["-O2", "-ffp-contract=off", "-fno-fast-math", "-DORACLE_ALT_TRIG"]}


def load_variants():
    so.build()
    for name, flags in VARIANTS.items():
        out = os.path.join(BUILD, f"libsrbd_oracle_{name}.so")
        subprocess.check_call(["gcc", *flags, "-fPIC", "-shared", "-o", out, os.path.join(ROOT, "oracle", "srbd_oracle.c"), "-lm"])
        so._LIBS[name] = C.CDLL(out)
    return [False] + list(VARIANTS)  # False = the strict checker build


def main():
    gpu = "--gpu" in sys.argv
    keys = load_variants()
    label = {False: "strict", "contract": "contract", "alt-trig": "alt-trig"}
    if gpu:
        import torch
        from quadruped_pympc_tamols_b200 import _cabi
    print(f"{'fixture':38s} {'N':>5s} | spread of the three CPU builds: cost rel, |dbest| | fixture (numpy stand-in) vs strict: cost rel, |dbest|"
          + (" | CUDA vs strict: cost rel, |dbest|" if gpu else ""))
    worst = dict(cpu_cost=0.0, cpu_best=0.0, gpu_cost=0.0, gpu_best=0.0)
    for path in sampling_fixtures():
        g = np.load(path)
        kw = fixture_kwargs(g)
        method, n = str(g["method"]), int(g["n"])
        cfg = so.make_cfg(**kw)
        eps = g["it0_eps"]
        freqs = g["it0_freqs"] if "it0_freqs" in g else None
        outs = {}
        for k in keys:
            outs[k] = so.solve(cfg, method, g["state"], g["reference"], g["contact_sequence"], g["best_in"], eps,
                               timing=g["timing"], freqs=freqs, fast=k)
        sc = max(rel_err(outs[a]["costs"], outs[b]["costs"]) for a in keys for b in keys if a is not b)
        sb = max(float(np.max(np.abs(outs[a]["best"] - outs[b]["best"]))) for a in keys for b in keys if a is not b)
        fc = rel_err(g["it0_costs"], outs[False]["costs"])
        fb = float(np.max(np.abs(g["it0_best_out"] - outs[False]["best"])))
        line = f"{os.path.basename(path)[9:-4]:38s} {n:5d} | {sc:9.2e} {sb:9.2e} | {fc:9.2e} {fb:9.2e}"
        worst["cpu_cost"], worst["cpu_best"] = max(worst["cpu_cost"], sc, fc), max(worst["cpu_best"], sb, fb)
        if gpu:
            h = _cabi.Handle(_cabi.make_config(sampling_method=method, num_parallel_computations=n,
                                               step_freq_available=tuple(g["step_freq_available"]), **kw))
            d_eps = torch.as_tensor(np.ascontiguousarray(eps, np.float32), device="cuda")
            d_fr = torch.as_tensor(np.ascontiguousarray(freqs, np.float32), device="cuda") if freqs is not None else None
            r = h.solve(g["state"], g["reference"], g["contact_sequence"], g["best_in"], 0, 0, timing=g["timing"],
                        nominal_freq=float(g["nominal_freq"]), optimize_swing=int(g["optimize_swing"]),
                        sigma_in=g["it0_sigma_in"] if method == "cem_mppi" else None, d_eps=d_eps.data_ptr(),
                        d_freqs=d_fr.data_ptr() if d_fr is not None else None)
            gc, gb = rel_err(r["costs"], outs[False]["costs"]), float(np.max(np.abs(r["best"] - outs[False]["best"])))
            worst["gpu_cost"], worst["gpu_best"] = max(worst["gpu_cost"], gc), max(worst["gpu_best"], gb)
            line += f" | {gc:9.2e} {gb:9.2e}  [{h.last_kernel}]"
            h.close()
        print(line)
    print(f"worst over the fixtures: CPU builds (incl. numpy stand-in) cost rel {worst['cpu_cost']:.2e}, |dbest| {worst['cpu_best']:.2e}"
          + (f"; CUDA vs strict cost rel {worst['gpu_cost']:.2e}, |dbest| {worst['gpu_best']:.2e}" if gpu else ""))


if __name__ == "__main__":
    main()
