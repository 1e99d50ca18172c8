"""CPU check of the error analysis behind the exact tree search (csrc/tree_kernels.cu header).

The device searches with an FP32 surrogate and decides in FP64 among the model points whose surrogate is <= thr2.
That is exact iff EVERY index attaining the reference's minimum (src/cpu.cc:16-22) passes the threshold, and iff no
box holding such a point can be skipped.  Here the FP32 arithmetic of the kernels is emulated with numpy float32
(an FMA = the exact product and sum in float64, rounded once more to float32) and both properties are asserted on
adversarial clouds.  The device rounds thr2 UP and box distances DOWN at every operation; this test uses the
real-valued formulas, i.e. a smaller threshold and a larger box distance than the device's: if the properties hold
here they hold there.
"""
import numpy as np
import pytest

F = np.float32
U = 2.0 ** -24
C_SQRT = float(F(1.0000002))        # the kernel's (1 + 3u) factor
C_2E = float(F(1.1920932e-7))       # the kernel's 2^-23 (1 + 3u)
C_THR = float(F(1.000001))          # the kernel's (1 + 16u) factor
C_RM = float(F(1.000001))


def fma32(a, b, c):
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(F)


def shadows(m, p):
    lo, hi = m.min(0), m.max(0)
    c0 = 0.5 * (lo + hi)
    h = np.maximum(hi - c0, c0 - lo)
    rm = np.sqrt((h * h).sum()) * (1.0 + 1e-9)                     # model_geo_kernel
    return (m - c0).astype(F), (p - c0).astype(F), rm


def surrogate(ph, mh):
    """d~2[j, k] with the operation order of surrogate_direct()."""
    dx = ph[:, None, 0] - mh[None, :, 0]
    dy = ph[:, None, 1] - mh[None, :, 1]
    dz = ph[:, None, 2] - mh[None, :, 2]
    return fma32(dz, dz, fma32(dy, dy, dx * dx))


def reference_values(p, m):
    d = p[:, None, :] - m[None, :, :]
    return np.sqrt((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2])


def threshold(best2, ph, rm):
    h = np.sqrt((ph.astype(np.float64) ** 2).sum(1))
    two_e = C_2E * (h + rm * C_RM) + 1e-30
    a = np.sqrt(best2.astype(np.float64)) * C_SQRT + two_e
    return a * a * C_THR


def cases():
    rng = np.random.default_rng(42)
    base = rng.uniform(-1, 1, (3000, 3))
    yield "uniform", base, rng.uniform(-1, 1, (300, 3))
    yield "scene = model + 1e-9", base, base[:300] + 1e-9
    yield "large common offset", base + 1000.0, base[:300] + 1000.0 + 1e-7
    yield "far scene", base, rng.uniform(-1, 1, (300, 3)) + np.array([40.0, -25.0, 60.0])
    yield "tiny scale", base * 1e-9, rng.uniform(-1, 1, (300, 3)) * 1e-9
    yield "huge scale", base * 1e9, rng.uniform(-1, 1, (300, 3)) * 1e9
    p = rng.uniform(-1, 1, (300, 3)) + 100.0                      # candidates 1e-12 apart (relative)
    d = rng.standard_normal((300, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    m = np.concatenate([p + 0.01 * d * (1 + 1e-12), p - 0.01 * d, p + 0.01 * d[:, [1, 2, 0]] * (1 + 3e-12)])
    yield "near ties", m[rng.permutation(len(m))], p
    yield "exact duplicates", np.concatenate([base[:1000], base[:1000][::-1]]), base[rng.permutation(1000)[:300]] - 1e-10
    flat = base.copy()
    flat[:, 2] = 0.25
    yield "planar model", flat, np.stack([np.linspace(-2, 2, 300), np.zeros(300), np.zeros(300)], 1)
    yield "fp32-exact grid", np.round(base * 64) / 64, np.round(rng.uniform(-1, 1, (300, 3)) * 64) / 64


@pytest.mark.parametrize("name,m,p", list(cases()), ids=[c[0] for c in cases()])
def test_every_reference_minimiser_passes_the_threshold(name, m, p):
    mh, ph, rm = shadows(m, p)
    d2 = surrogate(ph, mh)
    v = reference_values(p, m)
    best2 = d2.min(1)
    thr2 = threshold(best2, ph, rm)
    k0 = d2.argmin(1)
    v_k0 = v[np.arange(len(p)), k0]
    must_pass = v <= v_k0[:, None]          # superset of the reference's minimisers (what the proof bounds)
    assert must_pass[np.arange(len(p)), v.argmin(1)].all()
    worst = (d2.astype(np.float64) / thr2[:, None])[must_pass].max()
    assert worst <= 1.0, (name, worst)
    # the threshold is not vacuous: on average only a handful of points pass it
    if name in ("uniform", "far scene", "tiny scale", "huge scale"):
        assert (d2 <= thr2[:, None]).sum(1).mean() < 3.0


@pytest.mark.parametrize("name,m,p", list(cases()), ids=[c[0] for c in cases()])
def test_no_box_holding_a_candidate_can_be_skipped(name, m, p):
    mh, ph, _ = shadows(m, p)
    d2 = surrogate(ph, mh).astype(np.float64)
    for leaf in (8, 32):                                           # binary form / wide form
        n = len(mh) // leaf * leaf
        grp = mh[:n].reshape(-1, leaf, 3).astype(np.float64)
        lo, hi = grp.min(1), grp.max(1)
        q = ph.astype(np.float64)
        e = np.maximum(np.maximum(lo[None] - q[:, None], q[:, None] - hi[None]), 0.0)
        bd2 = (e * e).sum(2)                                       # real-valued: the device's is rounded down
        dmin_in_box = d2[:, :n].reshape(len(p), -1, leaf).min(2)
        # a box is skipped when bd2 > thr2 and a point passes when d~2 <= thr2: bd2 <= d~2 (1 + 6u) for every point of
        # the box is what (1 + 16u) in thr2 pays for
        tiny = 1e-44                                               # d~2 underflows before the real-valued bound does
        assert (bd2 <= dmin_in_box * (1 + 6 * U) + tiny).all(), (name, leaf)
