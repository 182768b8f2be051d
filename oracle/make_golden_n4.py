"""Generates tests/golden/stance_torque.npz and tests/golden/observable_rows.npz by EXECUTING the reference's own source
(SURVEY 8(f) N4).  Runs only where /root/reference exists (the build container).

* stance torques: the four statements tau.<leg> = -np.matmul(...) of interfaces/wb_interface.py (:368-372), cut out of the file as text (the surrounding method needs
  MuJoCo objects), de-indented and exec'd on LegsAttr containers of random Jacobians (3 x nv, nv = 18 like a floating-base
  quadruped) with the legs' joint-velocity index lists -- the statements themselves, not a restatement.
* dataset rows: the class MatLogger of simulation/simulation.py (:78-377), cut out of the module by its AST node (the
  module imports mujoco and gym_quadruped at the top) and exec'd with numpy, pathlib, LegsAttr and scipy's savemat in its
  namespace; record_step is fed the controller observables of quadruped_pympc_wrapper.py:205-243 for a few robots and
  steps, and the header / data table it would write to the .mat file is stored.
"""
import ast
import os
import pathlib
import sys
import textwrap

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("QPMPC_REFERENCE", "/root/reference")
LEGS = ("FL", "FR", "RL", "RR")


def reference_stance_torque_statements():
    src = open(os.path.join(REF, "quadruped_pympc", "interfaces", "wb_interface.py")).read().split("\n")
    first = [k for k, l in enumerate(src) if l.strip().startswith("tau.FL = -np.matmul(")][0]  # line 369 (1-based) at 6f49384
    lines = src[first:first + 4]
    assert lines[0].strip().startswith("tau.FL = -np.matmul(feet_jac.FL[:, legs_qvel_idx.FL].T, nmpc_GRFs.FL)"), lines[0]
    assert lines[3].strip().startswith("tau.RR = -np.matmul("), lines[3]
    return textwrap.dedent("\n".join(lines))


def reference_matlogger(LegsAttr):
    path = os.path.join(REF, "simulation", "simulation.py")
    tree = ast.parse(open(path).read())
    node = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "MatLogger"][0]
    from scipy.io import savemat
    ns = dict(np=np, pathlib=pathlib, LegsAttr=LegsAttr, STANDARD_LEG_ORDER=["FL", "FR", "RL", "RR"], savemat=savemat,
              SCIPY_AVAILABLE=True)
    exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
    return ns["MatLogger"]


def main():
    sys.path.insert(0, HERE)
    import jax_numpy_standin as standin
    LegsAttr = standin.LegsAttr
    out_dir = os.path.join(ROOT, "tests", "golden")
    rng = np.random.default_rng(368)

    # ---- stance torques
    code = compile(reference_stance_torque_statements(), "wb_interface.py:368-371", "exec")
    R, nv = 16, 18
    idx = {l: np.arange(6 + 3 * k, 9 + 3 * k) for k, l in enumerate(LEGS)}  # legs_qvel_idx of a floating-base quadruped
    jac_full = rng.normal(size=(R, 4, 3, nv))
    grf = (rng.normal(size=(R, 12)) * 10.0 ** rng.integers(-2, 3, size=(R, 12))).astype(np.float32)
    grf[0, 3:6] = 0.0  # a swing leg
    tau = np.zeros((R, 4, 3))
    for r in range(R):
        ns = dict(np=np, tau=LegsAttr(*[np.zeros(3) for _ in range(4)]),
                  feet_jac=LegsAttr(*[jac_full[r, k] for k in range(4)]),
                  legs_qvel_idx=LegsAttr(*[idx[l] for l in LEGS]),
                  nmpc_GRFs=LegsAttr(*[grf[r, 3 * k:3 * k + 3] for k in range(4)]))  # float32, like the jax array upstream
        exec(code, ns)
        tau[r] = np.stack([ns["tau"][l] for l in LEGS])
    jac_legs = np.stack([jac_full[:, k][:, :, idx[l]] for k, l in enumerate(LEGS)], axis=1)  # [R,4,3,3] = feet_jac[leg][:, idx]
    np.savez_compressed(os.path.join(out_dir, "stance_torque.npz"), feet_jac=jac_legs, grf=grf, tau=tau)
    print("stance_torque.npz:", tau.shape)

    # ---- dataset rows (MatLogger)
    MatLogger = reference_matlogger(LegsAttr)
    obs_names = ("ref_base_height", "ref_base_angles", "ref_feet_pos", "nmpc_GRFs", "nmpc_footholds", "swing_time",
                 "phase_signal", "lift_off_positions")
    state_names = ("base_pos", "base_lin_vel", "base_ori_euler_xyz", "base_ang_vel")
    Rr, T = 3, 6
    src = dict(time=np.arange(T) * 0.002,
               base_pos=rng.normal(size=(T, Rr, 3)), base_lin_vel=rng.normal(size=(T, Rr, 3)),
               base_ori_euler_xyz=rng.normal(size=(T, Rr, 3)), base_ang_vel=rng.normal(size=(T, Rr, 3)),
               ref_base_height=rng.normal(size=(T, Rr)), ref_base_angles=rng.normal(size=(T, Rr, 3)),
               ref_feet_pos=rng.normal(size=(T, Rr, 4, 3)),
               nmpc_GRFs=rng.normal(size=(T, Rr, 4, 3)).astype(np.float32),       # jax float32 upstream
               nmpc_footholds=rng.normal(size=(T, Rr, 4, 3)),
               swing_time=rng.uniform(0, 0.3, size=(T, Rr, 4)), phase_signal=rng.uniform(0, 1, size=(T, Rr, 4)),
               lift_off_positions=rng.normal(size=(T, Rr, 4, 3)))
    headers, tables = [], []
    for r in range(Rr):
        lg = MatLogger(list(state_names), obs_names, filepath=f"/tmp/_golden_matlogger_{r}.mat", write_every_n_steps=10 ** 9)
        for t in range(T):
            state = {k: src[k][t, r] for k in state_names}
            ctrl = dict(ref_base_height=float(src["ref_base_height"][t, r]), ref_base_angles=src["ref_base_angles"][t, r],
                        # quadruped_pympc_wrapper.py:213-219: the reference feet positions go in as (3, 1) columns
                        ref_feet_pos=LegsAttr(*[src["ref_feet_pos"][t, r, k].reshape(3, 1) for k in range(4)]),
                        nmpc_GRFs=LegsAttr(*[src["nmpc_GRFs"][t, r, k] for k in range(4)]),
                        nmpc_footholds=LegsAttr(*[src["nmpc_footholds"][t, r, k] for k in range(4)]),
                        swing_time=list(src["swing_time"][t, r]), phase_signal=list(src["phase_signal"][t, r]),
                        lift_off_positions=LegsAttr(*[src["lift_off_positions"][t, r, k] for k in range(4)]))
            lg.record_step(float(src["time"][t]), state, ctrl)
        headers.append(np.array(lg.header))
        tables.append(np.array(lg.all_data, dtype=np.float64))
    assert all(np.array_equal(h, headers[0]) for h in headers)
    np.savez_compressed(os.path.join(out_dir, "observable_rows.npz"), header=headers[0], data=np.stack(tables),
                        state_names=np.array(state_names), obs_names=np.array(obs_names), **{f"src_{k}": v for k, v in src.items()})
    print("observable_rows.npz: header", len(headers[0]), "data", np.stack(tables).shape)


if __name__ == "__main__":
    main()
