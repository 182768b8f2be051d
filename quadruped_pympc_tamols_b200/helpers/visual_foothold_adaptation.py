"""Drop-in for quadruped_pympc/helpers/visual_foothold_adaptation.py (class VisualFootholdAdaptation).

Same constructor, attributes, method names, positional argument order, in-place mutation of `reference_footholds`
and return values as the reference class (visual_foothold_adaptation.py:38-231).  The 'tamols' strategy -- the
reference's Python loop over every heightmap cell with cKDTree height queries (:164-228) -- runs as ONE CUDA launch
for all legs through the C ABI of include/tamols.h (tamols_score_host).  There is no CPU fallback for it: without the
shared library or an sm_100 GPU the constructor raises.

What stays on the host, as upstream: the `initialized` state machine (:59-72), the 'height' strategy (a single sensor
lookup per leg, :104-108) and the height-only fallback when no candidate is feasible (:223-228).  The 'vfa' strategy
needs the closed-source VFA module upstream (:6-18) and raises ImportError here exactly like the reference does when
that module is absent.
"""
import copy

import numpy as np

from .. import _tamols_cabi as tc
from .._config import resolve_config
from .legs_attr import LegsAttr


class VisualFootholdAdaptation:
    def __init__(self, legs_order, adaptation_strategy='height', config=None, device=0):
        self.footholds_adaptation = LegsAttr(
            FL=np.array([0, 0, 0]), FR=np.array([0, 0, 0]), RL=np.array([0, 0, 0]), RR=np.array([0, 0, 0])
        )
        self.footholds_constraints = LegsAttr(FL=None, FR=None, RL=None, RR=None)
        self.initialized = False
        self.adaptation_strategy = adaptation_strategy
        self._handle = None
        self._device = device
        if adaptation_strategy == 'vfa':
            raise ImportError("VFA strategy requested but VFA module could not be imported.")  # :50-51
        if adaptation_strategy == 'tamols':
            cfg = resolve_config(config)
            self.tamols_params = cfg.simulation_params.get('tamols_params', {})  # live dict, re-read on every call
            self.robot_name = cfg.robot
            self._handle = tc.Handle(tc.make_params(self.tamols_params, self.robot_name), device=device,
                                     max_jobs=max(4, len(legs_order)))
            self._params_seen = copy.deepcopy(self.tamols_params)  # what the handle currently holds
            self._pack = {}

    # ---- state machine, visual_foothold_adaptation.py:59-72
    def update_footholds_adaptation(self, update_footholds_adaptation):
        self.footholds_adaptation = update_footholds_adaptation
        self.initialized = True

    def reset(self):
        self.initialized = False

    def get_footholds_adapted(self, reference_footholds):
        if self.initialized == False:  # noqa: E712  (same test as upstream)
            self.footholds_adaptation = reference_footholds
            return reference_footholds, self.footholds_constraints
        return self.footholds_adaptation, self.footholds_constraints

    # ---- the hot path
    def compute_adaptation(self, legs_order, reference_footholds, hip_positions, heightmaps, forward_vel,
                           base_orientation, base_orientation_rate, gait_phases=None, base_position=None,
                           current_contact=None, current_feet_pos=None, phase_signal=None):
        """visual_foothold_adaptation.py:86-231.  `phase_signal` is accepted because the caller passes it
        (interfaces/wb_interface.py:238); like `gait_phases` it is unused by the 'tamols' and 'height' branches."""
        for leg_name in legs_order:
            if heightmaps[leg_name].data is None:
                return False

        if self.adaptation_strategy == 'height':
            for leg_name in legs_order:
                height_adjustment = heightmaps[leg_name].get_height(reference_footholds[leg_name])
                if height_adjustment is not None:
                    reference_footholds[leg_name][2] = height_adjustment

        elif self.adaptation_strategy == 'tamols':
            self.forward_vel = forward_vel
            self.base_position = base_position
            self.current_contact = current_contact if current_contact is not None else np.array([0, 0, 0, 0])
            self.current_feet_pos = current_feet_pos
            results = self._score_on_gpu(legs_order, reference_footholds, hip_positions, heightmaps)
            for k, leg_name in enumerate(legs_order):
                r = results[k]
                if int(r["best_index"]) >= 0:  # :209-222
                    best = np.array(r["best_xyz"], np.float64)
                    reference_footholds[leg_name] = best
                    self.footholds_constraints[leg_name] = [np.array(r["vertex1"], np.float64),
                                                            np.array(r["vertex2"], np.float64)]
                else:  # no feasible candidate: height-only fallback on the ORIGINAL sensor, :223-228
                    height_adjustment = heightmaps[leg_name].get_height(reference_footholds[leg_name])
                    if height_adjustment is not None:
                        reference_footholds[leg_name][2] = height_adjustment
        self.update_footholds_adaptation(reference_footholds)
        return True

    def _score_on_gpu(self, legs_order, reference_footholds, hip_positions, heightmaps):
        """Describe the legs (sensor array address + the per-leg vectors) and make the single C call that packs the
        job records and runs the one scoring launch (tamols_score_sensors)."""
        n = len(legs_order)
        # self.tamols_params is read with .get() on every call upstream, so a changed dict takes effect here too
        if self.tamols_params != self._params_seen:
            self._handle.set_params(tc.make_params(self.tamols_params, self.robot_name))
            self._params_seen = copy.deepcopy(self.tamols_params)
        legs = self._pack.get(n)
        if legs is None:
            legs = self._pack[n] = np.zeros(n, tc.SENSOR_LEG_DTYPE)
        vel = None if self.forward_vel is None else np.asarray(self.forward_vel, np.float64)
        keep, ptr, shape, flags = [], [], [], []
        stab = np.zeros((n, 2, 2))  # base_xy, diag_foot
        for k, name in enumerate(legs_order):
            d = np.ascontiguousarray(heightmaps[name].data, np.float64)  # [R, C, 1, 3]
            if d.ndim != 4 or d.shape[2] != 1 or d.shape[3] != 3:
                raise tc.TamolsError(f"heightmap data of leg {name} has shape {d.shape}, expected [rows, cols, 1, 3]")
            keep.append(d)
            ptr.append(d.ctypes.data)
            shape.append(d.shape[:2])
            f = tc.FLAG_HAS_VELOCITY if vel is not None else 0
            # _compute_stability_cost preconditions, :631-664 (leg_id is the position in legs_order)
            diag = tc.DIAGONAL.get(name)
            in_stance = k < len(self.current_contact) and self.current_contact[k] == 1
            if (self.base_position is not None and self.current_feet_pos is not None and diag is not None
                    and not in_stance and hasattr(self.current_feet_pos, diag)):
                if vel is None:
                    raise TypeError("forward_vel is None but the stability cost needs it (as upstream, :672)")
                f |= tc.FLAG_STABILITY
                stab[k, 0] = np.asarray(self.base_position, np.float64)[:2]
                stab[k, 1] = np.asarray(getattr(self.current_feet_pos, diag), np.float64)[:2]
            flags.append(f)
        # one assignment per record field for all legs (numpy fills structured arrays field by field: slow per item)
        legs["data"] = ptr
        legs["rows"], legs["cols"] = np.asarray(shape, np.int32).T
        legs["flags"] = flags
        legs["seed"] = [np.asarray(reference_footholds[name], np.float64) for name in legs_order]
        legs["hip"] = [np.asarray(hip_positions[name], np.float64) for name in legs_order]
        legs["vel"] = vel if vel is not None else 0.0
        legs["base_xy"], legs["diag_foot"] = stab[:, 0], stab[:, 1]
        return self._handle.score_sensors(legs)  # `keep` holds the sensor arrays until the call returns

    def close(self):
        if self._handle is not None:
            self._handle.close()
            self._handle = None
