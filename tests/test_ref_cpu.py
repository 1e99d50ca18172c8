"""The oracle restatement (oracle/icp_oracle.cc) against the REFERENCE'S OWN src/cpu.cc, compiled unmodified by
`make -C oracle _ref` (in the build container, where /root/reference exists) against the Eigen stand-in of
oracle/eigen_shim.  What this pins: the loop structure, operation order, tie rule, error definition, threshold
and Horn algebra of the reference's source text.  What it cannot pin: Eigen's own numerics (see the shim header).
Skipped where oracle/_ref/libref_cpu.so does not exist."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from oracle import oracle as O

SO = os.path.join(ROOT, "oracle", "_ref", "libref_cpu.so")
pytestmark = pytest.mark.skipif(not os.path.exists(SO), reason="oracle/_ref not built (needs /root/reference)")
dp = C.POINTER(C.c_double)


def _d(a):
    return a.ctypes.data_as(dp)


@pytest.fixture(scope="module")
def ref():
    L = C.CDLL(SO)
    L.ref_closest_matrix.argtypes = [C.c_long, C.c_long, dp, dp, dp]
    L.ref_err_compute.argtypes = [C.c_long, dp, C.c_double, dp, dp, dp]
    L.ref_err_compute.restype = C.c_double
    L.ref_err_compute_alignment.argtypes = [C.c_long, dp, C.c_double, dp, dp, dp]
    L.ref_err_compute_alignment.restype = C.c_double
    L.ref_find_alignment.argtypes = [C.c_long, dp, dp, dp, dp, dp]
    L.ref_find_alignment.restype = C.c_double
    return L


@pytest.mark.parametrize("mname,pname,n_scene", [("cow_ref", "cow_tr1", 2903), ("cow_ref", "cow_tr2", 600),
                                                  ("horse_ref", "horse_tr1", 256), ("bun000", "bun045", 256)])
def test_closest_matrix_same_columns_as_reference_source(ref, clouds, mname, pname, n_scene):
    m = clouds(mname)
    p = np.ascontiguousarray(clouds(pname)[:: max(1, len(clouds(pname)) // n_scene)][:n_scene])
    Y_ref = np.empty_like(p)
    ref.ref_closest_matrix(len(p), len(m), _d(p), _d(m), _d(Y_ref))
    for mode in (O.NN_LITERAL, O.NN_MULSQRT):
        idx, Y = O.closest_matrix(p, m, mode)
        assert np.array_equal(Y, Y_ref), mode          # columns of m, copied: exact
        assert np.array_equal(m[idx], Y_ref)


def test_closest_matrix_first_minimum_on_ties(ref):
    m = np.array([[1.0, 0, 0], [-1.0, 0, 0], [1.0, 0, 0], [0, 5.0, 0], [0, 5.0, 0]])
    p = np.array([[0.0, 0, 0], [0.9, 0, 0], [0, 4.0, 0]])
    Y_ref = np.empty_like(p)
    ref.ref_closest_matrix(3, 5, _d(p), _d(m), _d(Y_ref))
    idx, Y = O.closest_matrix(p, m, O.NN_LITERAL)
    assert idx.tolist() == [0, 0, 3] and np.array_equal(Y, Y_ref)


def test_find_alignment_and_residuals_match_reference_source(ref, clouds):
    m, p = clouds("cow_ref"), clouds("cow_tr1")
    idx, Y = O.closest_matrix(p, m, O.NN_MULSQRT)
    s = C.c_double()
    R, t = np.empty(9), np.empty(3)
    err_ref = ref.ref_find_alignment(len(p), _d(p), _d(Y), C.byref(s), _d(R), _d(t))
    for eig in (O.EIG_INTENDED, O.EIG_LITERAL):
        a = O.find_alignment(p, Y, eig)
        assert abs(a["s"] - s.value) <= 1e-13 * abs(s.value)
        assert np.allclose(a["R"].ravel(), R, rtol=0, atol=1e-13)
        assert np.allclose(a["t"], t, rtol=0, atol=1e-13)
        assert abs(a["err"] - err_ref) <= 1e-11 * abs(err_ref)
    e_align = ref.ref_err_compute_alignment(len(p), _d(p), s.value, _d(R), _d(t), _d(Y))
    assert abs(e_align - O.err_compute_alignment(p, s.value, R.reshape(3, 3), t, Y)) <= 1e-12 * abs(e_align)
    p1 = p.copy()
    e_ref = ref.ref_err_compute(len(p), _d(p1), s.value, _d(R), _d(t), _d(Y))
    e_or, p2 = O.err_compute(p, s.value, R.reshape(3, 3), t, Y)
    assert np.array_equal(p1, p2)                       # same operation order: sr = s*R, (sr0 x + sr1 y + sr2 z) + t
    assert e_ref == e_or


def test_full_loop_matches_reference_source(clouds, golden, tmp_path):
    """The reference's ICP::find_corresponding on cow_ref / cow_tr1: 7 iterations (the reference's only pinned
    datum, benchmark/callgrind.out.76685), the same printed error values and the same new_p as the oracle."""
    m, p = clouds("cow_ref"), clouds("cow_tr1")
    np.save(tmp_path / "m.npy", m)
    np.save(tmp_path / "p.npy", p)
    code = f"""
import ctypes as C, numpy as np
L = C.CDLL({SO!r}); dp = C.POINTER(C.c_double)
m = np.load({str(tmp_path / 'm.npy')!r}); p = np.load({str(tmp_path / 'p.npy')!r}); out = np.empty_like(p)
L.ref_find_corresponding.argtypes = [C.c_long, dp, C.c_long, dp, C.c_int, dp]
L.ref_find_corresponding(len(m), m.ctypes.data_as(dp), len(p), p.ctypes.data_as(dp), 10, out.ctypes.data_as(dp))
np.save({str(tmp_path / 'out.npy')!r}, out)
"""
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stderr.splitlines() if l.startswith("[ICP] iteration number")]
    assert len(lines) == 7
    printed = np.array([float(l.split("error value = ")[1]) for l in lines])
    g = golden("cow_tr1")
    ro = O.icp_run(m, p, 10, nn_mode=O.NN_LITERAL, eig_mode=O.EIG_LITERAL)
    assert ro["iters"] == 7 == int(g["iters"])
    assert np.allclose(printed, ro["err"], rtol=5e-6, atol=0)        # the reference prints 6 significant digits
    out = np.load(tmp_path / "out.npy")
    assert np.abs(out - ro["new_p"]).max() < 1e-12
    assert np.abs(out - g["new_p"]).max() < 1e-12
