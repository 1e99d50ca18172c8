"""ctypes binding of oracle/liboracle_icp.so — the CPU restatement of the reference's src/cpu.cc.

TEST INFRASTRUCTURE ONLY (see icp_oracle.cc header): imported by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs, never by the product path.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


def build(force=False):
    so = os.path.join(_HERE, "liboracle_icp.so")
    src = os.path.join(_HERE, "icp_oracle.cc")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.oracle_num_threads.restype = C.c_int
        L.oracle_load_matrix.restype = C.c_long
        L.oracle_load_matrix.argtypes = [C.c_char_p, _dp, C.c_long]
        L.oracle_write_matrix.argtypes = [C.c_char_p, _dp, C.c_long]
        L.oracle_closest_matrix.restype = None
        L.oracle_closest_matrix.argtypes = [C.c_long, C.c_long, _dp, _dp, C.c_int, C.c_int, _ip, _dp]
        L.oracle_err_compute.restype = C.c_double
        L.oracle_err_compute.argtypes = [C.c_long, _dp, C.c_double, _dp, _dp, _dp]
        L.oracle_err_compute_alignment.restype = C.c_double
        L.oracle_err_compute_alignment.argtypes = [C.c_long, _dp, C.c_double, _dp, _dp, _dp]
        L.oracle_centroid.restype = None
        L.oracle_centroid.argtypes = [C.c_long, _dp, _dp, _dp, _dp, _dp, _dp]
        L.oracle_horn_matrix.restype = None
        L.oracle_horn_matrix.argtypes = [_dp, _dp]
        L.oracle_horn_quaternion.restype = C.c_int
        L.oracle_horn_quaternion.argtypes = [_dp, C.c_int, _dp, _dp]
        L.oracle_quat_to_rot.restype = None
        L.oracle_quat_to_rot.argtypes = [_dp, _dp]
        L.oracle_find_alignment.restype = C.c_double
        L.oracle_find_alignment.argtypes = [C.c_long, _dp, _dp, C.c_int, _dp, _dp, _dp, _dp, C.POINTER(C.c_int)]
        L.oracle_icp_run.restype = C.c_int
        L.oracle_icp_run.argtypes = [C.c_long, _dp, C.c_long, _dp, C.c_int, C.c_double, C.c_int, C.c_int,
                                     C.c_int, C.c_int, C.c_int, _dp, _ip, _dp]
        _LIB = L
    return _LIB


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _i(a):
    return a.ctypes.data_as(_ip) if a is not None else None


def _pts(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    assert a.ndim == 2 and a.shape[1] == 3, "clouds are (n, 3) float64 (the reference's 3 x n col-major)"
    return a


NN_LITERAL, NN_MULSQRT, NN_MUL = 0, 1, 2
EIG_INTENDED, EIG_LITERAL = 0, 1


def num_threads():
    return lib().oracle_num_threads()


def load_matrix(path):
    """src/load.cc:3-33 -> (n, 3) float64; raises FileNotFoundError where the reference exit(2)s."""
    n = lib().oracle_load_matrix(os.fsencode(path), None, 0)
    if n == -2:
        raise FileNotFoundError(path)
    out = np.zeros((max(n, 0), 3))
    lib().oracle_load_matrix(os.fsencode(path), _d(out), n)
    return out


def write_matrix(path, pts):
    pts = _pts(pts)
    return lib().oracle_write_matrix(os.fsencode(path), _d(pts), len(pts))


def closest_matrix(p, m, nn_mode=NN_MULSQRT, nthreads=0):
    """src/cpu.cc:5-27 -> (idx int32 (np,), Y (np,3))."""
    p, m = _pts(p), _pts(m)
    idx = np.empty(len(p), np.int32)
    Y = np.empty_like(p)
    lib().oracle_closest_matrix(len(p), len(m), _d(p), _d(m), nn_mode, nthreads, _i(idx), _d(Y))
    return idx, Y


def err_compute(p, s, R, t, Y):
    """src/cpu.cc:29-40 -> (err, new_p).  Input p is not modified (a copy is transformed)."""
    p = _pts(p).copy()
    R = np.ascontiguousarray(R, np.float64); t = np.ascontiguousarray(t, np.float64)
    e = lib().oracle_err_compute(len(p), _d(p), float(s), _d(R), _d(t), _d(_pts(Y)))
    return e, p


def err_compute_alignment(p, s, R, t, y):
    p = _pts(p)
    R = np.ascontiguousarray(R, np.float64); t = np.ascontiguousarray(t, np.float64)
    return lib().oracle_err_compute_alignment(len(p), _d(p), float(s), _d(R), _d(t), _d(_pts(y)))


def centroid(p, y, centred=False):
    p, y = _pts(p), _pts(y)
    mu_p, mu_y = np.empty(3), np.empty(3)
    pp = np.empty_like(p) if centred else None
    yp = np.empty_like(y) if centred else None
    lib().oracle_centroid(len(p), _d(p), _d(y), _d(mu_p), _d(mu_y), _d(pp), _d(yp))
    return (mu_p, mu_y, pp, yp) if centred else (mu_p, mu_y)


def horn(S, eig_mode=EIG_INTENDED):
    """S (3,3) cross-covariance -> (q (4,), R (3,3), eigenvalues (4,), picked column)."""
    S = np.ascontiguousarray(S, np.float64)
    N = np.empty(16); q = np.empty(4); ev = np.empty(4); R = np.empty(9)
    lib().oracle_horn_matrix(_d(S), _d(N))
    k = lib().oracle_horn_quaternion(_d(N), eig_mode, _d(q), _d(ev))
    lib().oracle_quat_to_rot(_d(q), _d(R))
    return q, R.reshape(3, 3), ev, k


def find_alignment(p, y, eig_mode=EIG_INTENDED):
    """src/cpu.cc:105-175 -> dict(s, R (3,3), t (3,), err, S (3,3), picked)."""
    p, y = _pts(p), _pts(y)
    s = C.c_double(); R = np.empty(9); t = np.empty(3); S = np.empty(9); k = C.c_int()
    err = lib().oracle_find_alignment(len(p), _d(p), _d(y), eig_mode, C.byref(s), _d(R), _d(t), _d(S), C.byref(k))
    return dict(s=s.value, R=R.reshape(3, 3), t=t, err=err, S=S.reshape(3, 3), picked=k.value)


def icp_run(m, p, max_iter, threshold=1e-5, nn_mode=NN_MULSQRT, eig_mode=EIG_INTENDED, allow_unequal=False,
            nthreads=0, log=False, keep_idx=False):
    """src/cpu.cc:55-79 -> dict(iters, err (iters,), idx (iters, np) | None, s, R, t per iteration, new_p).

    iters < 0 mirrors the reference's exit(-1) conditions (-1 unequal sizes, -4 fewer than 4 points).
    """
    m, p = _pts(m), _pts(p).copy()
    err = np.zeros(max(max_iter, 1)); srt = np.zeros((max(max_iter, 1), 13))
    idx = np.zeros((max(max_iter, 1), len(p)), np.int32) if keep_idx else None
    it = lib().oracle_icp_run(len(m), _d(m), len(p), _d(p), max_iter, threshold, nn_mode, eig_mode,
                              int(allow_unequal), nthreads, int(log), _d(err), _i(idx), _d(srt))
    n = max(it, 0)
    return dict(iters=it, err=err[:n], idx=None if idx is None else idx[:n], s=srt[:n, 0],
                R=srt[:n, 1:10].reshape(n, 3, 3), t=srt[:n, 10:13], new_p=p)
