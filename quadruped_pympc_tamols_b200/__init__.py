"""B200-native (sm_100a) drop-in for the sampling-MPC and TAMOLS hot path of Quadruped-PyMPC-TAMOLS.

Only the hot path lives here: csrc/ (CUDA kernels + C ABI, see include/*.h), _cabi.py (ctypes binding) and the
host-side mirrors of the reference's two classes (controllers/sampling/*, helpers/visual_foothold_adaptation.py).
"""
__version__ = "0.1.0"
