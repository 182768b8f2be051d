"""CPU: the arithmetic of csrc/srbd.cuh::sincos_fast (three-term Cody-Waite reduction + minimax polynomials, the
branch-free replacement of the library sincosf in the rollout step) restated in numpy float32 with explicitly rounded
fused multiply-adds, against float64 sin/cos: it must stay in the library's accuracy class (<= 2 ulp, 1e-7 absolute),
because the temperature-1 softmin amplifies sin/cos errors of 5e-7 beyond the 1e-3 bar."""
import numpy as np

f32 = np.float32


def _fma(a, b, c):
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def sincos_fast(a):
    a = a.astype(f32)
    b = np.where(np.abs(a) < f32(105615.0), a, f32(0.0)).astype(f32)
    q = np.rint((b * f32(0.636619772)).astype(f32)).astype(np.int32)
    k = q.astype(f32)
    C = lambda v: np.full_like(b, v)  # noqa: E731
    r = _fma(k, C(-1.5707962513e+0), b)
    r = _fma(k, C(-7.5497894159e-8), r)
    r = _fma(k, C(-5.3903029535e-15), r)
    r2 = (r * r).astype(f32)
    c = _fma(r2, C(2.443315711809948e-5), C(-1.388731625493765e-3))
    c = _fma(c, r2, C(4.166664568298827e-2))
    c = _fma((c * r2).astype(f32), r2, _fma(C(-0.5), r2, C(1.0)))
    s = _fma(r2, C(-1.9515295891e-4), C(8.3321608736e-3))
    s = _fma(s, r2, C(-1.6666654611e-1))
    s = _fma((s * r2).astype(f32), r, r)
    odd = (q & 1) == 1
    ss, cc = np.where(odd, c, s), np.where(odd, s, c)
    return np.where((q & 2) != 0, -ss, ss), np.where(((q + 1) & 2) != 0, -cc, cc)


def test_accuracy_class_of_the_library_sincosf():
    rng = np.random.default_rng(0)
    for scale in (0.05, 1.0, 10.0, 1.0e3, 1.0e5):
        a = rng.uniform(-scale, scale, 400_000).astype(f32)
        s, c = sincos_fast(a)
        rs, rc = np.sin(a.astype(np.float64)), np.cos(a.astype(np.float64))
        assert np.max(np.abs(s - rs)) < 1.0e-7 and np.max(np.abs(c - rc)) < 1.0e-7
        assert np.max(np.abs(s - rs) / np.spacing(np.abs(rs).astype(f32))) < 2.0
        assert np.max(np.abs(c - rc) / np.spacing(np.abs(rc).astype(f32))) < 2.0


def test_quadrants_and_out_of_range_inputs():
    a = np.array([0.0, np.pi / 2, np.pi, 3 * np.pi / 2, -np.pi / 2, 2 * np.pi, 0.7853982, -2.3561945], f32)
    s, c = sincos_fast(a)
    assert np.allclose(s, np.sin(a.astype(np.float64)), atol=1e-7) and np.allclose(c, np.cos(a.astype(np.float64)), atol=1e-7)
    s, c = sincos_fast(np.array([1.0e6, -3.0e38, np.inf, np.nan, 105615.0], f32))
    assert np.all(s == 0.0) and np.all(c == 1.0)  # evaluated at 0: the rollout is lost through its angle cost term
