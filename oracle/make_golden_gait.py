"""Generates tests/golden/gait_sequences.npz from the reference's OWN PeriodicGaitGenerator
(/root/reference/quadruped_pympc/helpers/periodic_gait_generator.py, unmodified, imported under the gym_quadruped /
mujoco.viewer stubs of oracle/jax_numpy_standin.py).  Runs only where /root/reference exists (the build container).

Each case: a clock of some gait advanced by a random number of run(dt, f) calls (so the phases are generic floats),
optionally put into the `init` hold state, then compute_contact_sequence with a uniform or non-uniform dt list; a few
more run() steps record the stepping semantics (state carried across calls)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("QPMPC_REFERENCE", "/root/reference")


def main():
    sys.path.insert(0, HERE)
    import jax_numpy_standin as standin
    standin.install_robot_stubs()
    for k in [k for k in sys.modules if k == "quadruped_pympc" or k.startswith("quadruped_pympc.")]:
        del sys.modules[k]
    sys.path.insert(0, REF)
    from quadruped_pympc.helpers.periodic_gait_generator import PeriodicGaitGenerator
    import quadruped_pympc.helpers.periodic_gait_generator as mod
    assert os.path.abspath(mod.__file__).startswith(os.path.abspath(REF))

    rng = np.random.default_rng(2024)
    cases = dict(phase=[], init=[], offset=[], freq=[], duty=[], dts=[], lengths=[], horizon=[], seq=[], gait=[],
                 run_dt=[], run_contacts=[], run_phase=[], run_init=[])
    H = 12
    for gait in range(7):
        for rep in range(6):
            duty = float(rng.choice([0.5, 0.6, 0.65, 0.8, 0.9]))
            freq = float(rng.choice([0.5, 1.3, 1.4, 2.0, 2.4]))
            g = PeriodicGaitGenerator(duty_factor=duty, step_freq=freq, gait_type=gait, horizon=H)
            for _ in range(int(rng.integers(0, 40))):
                g.run(float(rng.choice([0.002, 0.01, 0.02])), freq)
            if rep in (1, 4):  # hold state: legs wait for their phase offset before the first lift-off (:58-65)
                g.set_phase_signal(np.array(g.phase_signal) * 0.5, init=[True, False, True, True])
            if rep % 2 == 0:
                dts, lengths = [0.02], [H]
            else:  # non-uniform discretisation as wb_interface passes it: fine steps first, then coarse
                dts, lengths = [0.01, 0.02], [int(rng.integers(2, 6)), H]
            cases["phase"].append(np.array(g._phase_signal, np.float64))
            cases["init"].append(np.array(g._init, bool))
            cases["offset"].append(np.array(g.phase_offset, np.float64))
            cases["freq"].append(freq)
            cases["duty"].append(duty)
            cases["dts"].append(np.array(dts + [0.0] * (2 - len(dts))))
            cases["lengths"].append(np.array(lengths + [0] * (2 - len(lengths))))
            cases["horizon"].append(H)
            cases["gait"].append(gait)
            seq = g.compute_contact_sequence(contact_sequence_dts=dts, contact_sequence_lenghts=lengths)
            assert np.array_equal(np.array(g._phase_signal), cases["phase"][-1])  # state restored
            cases["seq"].append(np.array(seq))
            rdt = float(rng.choice([0.002, 0.01]))
            cs = []
            for _ in range(8):
                cs.append(np.array(g.run(rdt, freq)))
            cases["run_dt"].append(rdt)
            cases["run_contacts"].append(np.array(cs))
            cases["run_phase"].append(np.array(g._phase_signal, np.float64))
            cases["run_init"].append(np.array(g._init, bool))
    out = os.path.join(os.path.dirname(HERE), "tests", "golden", "gait_sequences.npz")
    np.savez_compressed(out, **{k: np.array(v) for k, v in cases.items()})
    print("wrote", out, len(cases["seq"]), "cases")


if __name__ == "__main__":
    main()
