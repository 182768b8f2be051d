"""LegsAttr: the 4-field container the reference passes across the TAMOLS boundary.

The reference takes it from the third-party package gym_quadruped (gym_quadruped.utils.quadruped_utils.LegsAttr,
imported at quadruped_pympc/helpers/visual_foothold_adaptation.py:3).  When that package is installed its class is
re-exported unchanged; otherwise a minimal stand-in with the surface the TAMOLS path uses (attribute access, item
access by leg name, assignment) is provided so the drop-in class can be exercised without the simulator stack.
"""
try:  # pragma: no cover - depends on the user's environment
    from gym_quadruped.utils.quadruped_utils import LegsAttr  # noqa: F401
except Exception:  # noqa: BLE001

    class LegsAttr:
        order = ("FL", "FR", "RL", "RR")

        def __init__(self, FL, FR, RL, RR, **_):
            self.FL, self.FR, self.RL, self.RR = FL, FR, RL, RR

        def __getitem__(self, key):
            return getattr(self, key if isinstance(key, str) else self.order[key])

        def __setitem__(self, key, value):
            setattr(self, key if isinstance(key, str) else self.order[key], value)

        def __iter__(self):
            return iter((self.FL, self.FR, self.RL, self.RR))

        def to_dict(self):
            return {k: getattr(self, k) for k in self.order}

        def __repr__(self):
            return f"LegsAttr(FL={self.FL}, FR={self.FR}, RL={self.RL}, RR={self.RR})"
