"""numpy-backed stand-in for the slice of the jax API the reference's sampling controller uses.

TEST INFRASTRUCTURE ONLY.  jax/jaxlib are not installable in this container, so the reference's
`Sampling_MPC` (controllers/sampling/centroidal_nmpc_jax*.py), `Centroidal_Model_JAX` and
`PeriodicGaitGeneratorJax` cannot run on XLA here.  `install()` registers fake `jax`, `jax.numpy`,
`jax.random`, `jax.lax` modules (plus the `gym_quadruped` / `mujoco.viewer` stubs config.py needs) in
`sys.modules`, so that the reference's UNMODIFIED Python source executes with numpy arithmetic that
follows jax's dtype rules (float32 / int32 everywhere, weak Python scalars, clamped/wrapped scalar
indexing, functional `.at[].set`).  oracle/make_golden_sampling.py uses it to produce the fixtures
that pin oracle/srbd_oracle.c.  Random draws cannot reproduce jax's threefry streams; they are
deterministic numpy draws keyed by the jax key, and the fixtures store the drawn arrays explicitly.
"""
import sys
import types

import numpy as np

_F64, _I64 = np.dtype(np.float64), np.dtype(np.int64)


def _narrow(a):
    a = np.asarray(a)
    if a.dtype == _F64:
        a = a.astype(np.float32)
    elif a.dtype == _I64:
        a = a.astype(np.int32)
    return a


def _wrap(a):
    if isinstance(a, tuple):
        return tuple(_wrap(x) for x in a)
    return _narrow(a).view(JArray)


def _base(x):
    """Operand for a numpy call: arrays narrowed to jax's default widths, Python scalars left weak."""
    if isinstance(x, (bool, int, float)):
        return x
    if isinstance(x, (list, tuple)):
        return _narrow(np.array([_base(v) for v in x]))
    return _narrow(np.asarray(x))


def _promote(args):
    """jax promotes int32 (+) float32 -> float32 where numpy would give float64."""
    arrs = [a for a in args if isinstance(a, np.ndarray)]
    if any(a.dtype.kind == "f" for a in arrs):
        return [a.astype(np.float32) if isinstance(a, np.ndarray) and a.dtype.kind in "iu" else a for a in args]
    return args


class _At:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtIdx(self.arr, idx)


class _AtIdx:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def set(self, value):
        out = np.array(self.arr, copy=True).view(np.ndarray)
        out[_clean_index(self.idx, out.shape)] = np.asarray(_base(value)).astype(out.dtype)
        return _wrap(out)

    def get(self):
        return self.arr[self.idx]


def _clean_index(idx, shape):
    if isinstance(idx, tuple):
        return tuple(_clean_index(i, (shape[k],) if k < len(shape) else ()) for k, i in enumerate(idx))
    if isinstance(idx, np.ndarray) and idx.ndim == 0 and idx.dtype.kind in "iu":
        idx = int(idx)
    if isinstance(idx, (int, np.integer)) and len(shape) >= 1:
        i, n = int(idx), shape[0]
        if i < 0:
            i += n
        return min(max(i, 0), n - 1)  # jax clamps out-of-range scalar gathers
    return idx


class JArray(np.ndarray):
    __array_priority__ = 1000

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kwargs):
        args = _promote([_base(x) for x in inputs])
        res = getattr(ufunc, method)(*args, **kwargs)
        return _wrap(res)

    def __array_function__(self, func, types_, args, kwargs):
        def strip(x):
            if isinstance(x, JArray):
                return x.view(np.ndarray)
            if isinstance(x, (list, tuple)):
                return type(x)(strip(v) for v in x)
            return x
        res = func(*strip(args), **{k: strip(v) for k, v in kwargs.items()})
        if isinstance(res, np.ndarray):
            return _wrap(res)
        if isinstance(res, (list, tuple)) and all(isinstance(r, np.ndarray) for r in res):
            return type(res)(_wrap(r) for r in res)
        return res

    def __getitem__(self, idx):
        base = self.view(np.ndarray)
        return _wrap(np.ndarray.__getitem__(base, _clean_index(idx, base.shape)))

    def __iter__(self):
        for i in range(self.shape[0]):
            yield self[i]

    @property
    def at(self):
        return _At(self)

    def take(self, idx):
        return self[idx]


def _make_jnp():
    jnp = types.ModuleType("jax.numpy")
    jnp.ndarray = JArray
    jnp.newaxis = None
    jnp.float32 = np.float32
    jnp.int16 = lambda x: _wrap(np.int16(np.asarray(x)))

    def array(x, dtype=None):
        a = _base(x) if not isinstance(x, np.ndarray) else _narrow(x)
        a = np.asarray(a)
        if dtype is not None:
            a = a.astype(dtype)
        return _wrap(a)

    jnp.array = array
    jnp.asarray = array
    jnp.zeros = lambda shape, dtype=np.float32: _wrap(np.zeros(shape, dtype))
    jnp.ones = lambda shape, dtype=np.float32: _wrap(np.ones(shape, dtype))
    jnp.identity = lambda n, dtype=np.float32: _wrap(np.identity(n, dtype))
    jnp.eye = lambda n, dtype=np.float32: _wrap(np.eye(n, dtype=dtype))
    jnp.arange = lambda *a: _wrap(np.arange(*a))
    jnp.linspace = lambda a, b, n: _wrap(np.linspace(a, b, n))

    def tile(x, reps):
        return _wrap(np.tile(_base(x), int(reps) if not isinstance(reps, tuple) else reps))

    jnp.tile = tile
    jnp.concatenate = lambda xs, axis=0: _wrap(np.concatenate([np.atleast_1d(_base(x)) for x in xs], axis=axis))
    jnp.split = lambda x, n: [_wrap(p) for p in np.split(_base(x), n)]

    def dot(a, b):
        a, b = _promote([_base(a), _base(b)])
        return _wrap(np.dot(a, b))

    jnp.dot = dot
    for name in ("sin", "cos", "exp", "sqrt", "isnan", "isinf", "abs"):
        setattr(jnp, name, (lambda f: lambda x: _wrap(f(_base(x))))(getattr(np, name)))
    jnp.max = lambda x: _wrap(np.max(_base(x)))
    jnp.sum = lambda x, axis=None: _wrap(np.sum(_base(x), axis=axis))

    def where(c, a, b):
        a, b = _promote([_base(a), _base(b)])
        return _wrap(np.where(_base(c), a, b))

    jnp.where = where
    jnp.nanargmin = lambda x: _wrap(np.nanargmin(_base(x)))
    jnp.argsort = lambda x: _wrap(np.argsort(_base(x), kind="stable"))
    jnp.cov = lambda x, rowvar=True: _wrap(np.cov(_base(x).astype(np.float64), rowvar=rowvar))
    jnp.diag = lambda x: _wrap(np.diag(_base(x)))
    return jnp


DRAWS = []  # (kind, array) of every random draw, in call order, for the fixture writer


def _rng(key):
    return np.random.default_rng([int(v) for v in np.asarray(key).ravel()])


def _make_random():
    r = types.ModuleType("jax.random")
    r.PRNGKey = lambda seed: _wrap(np.array([0, seed], np.uint32))

    def split(key, num=2):
        g = _rng(key)
        return _wrap(g.integers(0, 2**32, size=(num, 2), dtype=np.uint32))

    def normal(key, shape):
        a = _rng(key).standard_normal(shape).astype(np.float32)
        DRAWS.append(("normal", a))
        return _wrap(a)

    def uniform(key, minval=0.0, maxval=1.0, shape=()):
        a = _rng(key).uniform(minval, maxval, shape).astype(np.float32)
        DRAWS.append(("uniform", a))
        return _wrap(a)

    def choice(key, a, shape=()):
        a = np.asarray(_base(a))
        out = a[_rng(key).integers(0, a.shape[0], size=shape)]
        DRAWS.append(("choice", np.array(out)))
        return _wrap(out)

    r.split, r.normal, r.uniform, r.choice = split, normal, uniform, choice
    return r


def _jit(fun, device=None, **_):
    def wrapped(*args, **kwargs):
        conv = lambda x: _wrap(x) if isinstance(x, (np.ndarray, list)) else x
        return fun(*[conv(a) for a in args], **{k: conv(v) for k, v in kwargs.items()})
    return wrapped


def _vmap(fun, in_axes=0, out_axes=0):
    def mapped(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = next(np.asarray(a).shape[0] for a, ax in zip(args, axes) if ax is not None)
        outs = [fun(*[(_wrap(np.asarray(a))[i] if ax is not None else a) for a, ax in zip(args, axes)])
                for i in range(n)]
        return _wrap(np.stack([np.asarray(o) for o in outs]))
    return mapped


def _fori_loop(lo, hi, body, init):
    carry = init
    for n in range(int(lo), int(hi)):
        carry = body(n, carry)
    return carry


def install(robot_hip_height=0.33):
    """Register the stand-in modules.  Idempotent."""
    if "jax" in sys.modules and getattr(sys.modules["jax"], "_is_standin", False):
        return sys.modules["jax"]
    jax = types.ModuleType("jax")
    jax._is_standin = True
    jnp, rnd = _make_jnp(), _make_random()
    lax = types.ModuleType("jax.lax")
    lax.fori_loop = _fori_loop
    jax.numpy, jax.random, jax.lax = jnp, rnd, lax
    jax.jit, jax.vmap = _jit, _vmap
    jax.devices = lambda kind=None: (_ for _ in ()).throw(RuntimeError("no gpu")) if kind == "gpu" else ["cpu0"]
    jax.config = types.SimpleNamespace(update=lambda *a, **k: None)
    jax.debug = types.SimpleNamespace(breakpoint=lambda: None)
    sys.modules.update({"jax": jax, "jax.numpy": jnp, "jax.random": rnd, "jax.lax": lax})
    install_robot_stubs(robot_hip_height)
    return jax


class LegsAttr:
    """Stand-in for gym_quadruped.utils.quadruped_utils.LegsAttr (usage-defined contract: a mutable
    4-field container addressable by attribute and by leg name)."""

    def __init__(self, FL=None, FR=None, RL=None, RR=None):
        self.FL, self.FR, self.RL, self.RR = FL, FR, RL, RR

    def __getitem__(self, k):
        return getattr(self, k) if isinstance(k, str) else getattr(self, ("FL", "FR", "RL", "RR")[k])

    def __setitem__(self, k, v):
        setattr(self, k if isinstance(k, str) else ("FL", "FR", "RL", "RR")[k], v)

    def __add__(self, other):  # gym_quadruped's LegsAttr adds an array (or another LegsAttr) to every leg
        o = (lambda k: getattr(other, k)) if isinstance(other, LegsAttr) else (lambda k: other)
        return LegsAttr(*[getattr(self, k) + o(k) for k in ("FL", "FR", "RL", "RR")])

    __iadd__ = __add__


def install_robot_stubs(hip_height=0.33):
    """gym_quadruped / mujoco.viewer stubs needed to import quadruped_pympc.config (config.py:11-16,
    helpers/quadruped_utils.py:8)."""
    if "gym_quadruped" in sys.modules:
        return
    gq = types.ModuleType("gym_quadruped")
    utils = types.ModuleType("gym_quadruped.utils")
    qu = types.ModuleType("gym_quadruped.utils.quadruped_utils")
    qu.LegsAttr = LegsAttr
    rc = types.ModuleType("gym_quadruped.robot_cfgs")

    class RobotConfig:
        def __init__(self):
            self.leg_joints = LegsAttr(*[[f"{l}_{j}" for j in ("hip", "thigh", "calf")] for l in ("FL", "FR", "RL", "RR")])
            self.feet_geom_names = LegsAttr("FL", "FR", "RL", "RR")
            self.qpos0_js = np.zeros(12)
            self.hip_height = hip_height

    rc.RobotConfig = RobotConfig
    rc.get_robot_config = lambda robot_name: RobotConfig()
    mj = types.ModuleType("mujoco")
    mjv = types.ModuleType("mujoco.viewer")
    mjv.Handle = object
    mj.viewer = mjv
    gq.utils, gq.robot_cfgs, utils.quadruped_utils = utils, rc, qu
    sys.modules.update({"gym_quadruped": gq, "gym_quadruped.utils": utils, "gym_quadruped.utils.quadruped_utils": qu,
                        "gym_quadruped.robot_cfgs": rc, "mujoco": mj, "mujoco.viewer": mjv})
