"""GPU test of icp_find_alignment_noalloc (include/icp_b200.h): the no-allocation variant of the fused moments + Horn
solve, called through the C-ABI on device buffers, against the existing entry point icp_find_alignment_host on the same
pairs - two calls in a row with different data (the entry point's scratch is one shared device array that keeps the last
call's contents).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import synthetic_pair

pytestmark = pytest.mark.gpu

RTOL = 1e-5      # the tolerance of tests/test_gpu_parity.py for floating-point results (the two paths agree to ~1e-12)


def _noalloc(L, torch, y, p):
    """-> 20 doubles (s, R[9], t[3], mu_p[3], mu_y[3], 0).  Device buffers are torch's; the entry point allocates nothing."""
    dy = torch.from_numpy(np.ascontiguousarray(y)).cuda()
    dp = torch.from_numpy(np.ascontiguousarray(p)).cuda()
    out = torch.full((20,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.icp_find_alignment_noalloc(C.c_void_p(dy.data_ptr()), C.c_void_p(dp.data_ptr()), len(p), C.c_void_p(out.data_ptr()),
                                      C.c_void_p(stream))
    assert rc == 0, L.icp_last_error()
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _check(icp_b200, got, y, p):
    ref = icp_b200.find_alignment(p, y)                  # icp_find_alignment_host: s, R, t (and the residual) on the same pairs
    s, R, t = got[0], got[1:10].reshape(3, 3), got[10:13]
    assert abs(s - ref["s"]) <= RTOL * abs(ref["s"])
    assert np.abs(R - ref["R"]).max() <= RTOL
    assert np.abs(t - ref["t"]).max() <= RTOL * max(1.0, np.abs(ref["t"]).max())
    assert np.abs(got[13:16] - p.mean(0)).max() <= 1e-9 * max(1.0, np.abs(p).max())
    assert np.abs(got[16:19] - y.mean(0)).max() <= 1e-9 * max(1.0, np.abs(y).max())
    # far tighter than the gate in practice: both paths add FP64 partial sums, only the pivot and the order differ
    assert abs(s - ref["s"]) <= 1e-9 and np.abs(R - ref["R"]).max() <= 1e-9


def test_noalloc_find_alignment_matches_the_existing_entry_point_twice():
    import torch
    import icp_b200
    L = icp_b200.load_library()
    torch.cuda.set_device(0)
    # call 1: 200 000 noisy pairs far from the origin (782 blocks' worth: the block cap is exercised)
    m1, p1, perm1 = synthetic_pair(200000, seed=51, angle_deg=7.0, shift=(40.0, -25.0, 60.0))
    y1 = m1[perm1] * 1.3                                   # a scale != 1, so that s is exercised
    got1 = _noalloc(L, torch, y1, p1)
    _check(icp_b200, got1, y1, p1)
    # call 2, right after, other size and data: must not see anything of call 1 in the shared scratch
    m2, p2, perm2 = synthetic_pair(3001, seed=52, angle_deg=2.0, noise=0.0)
    y2 = m2[perm2]
    got2 = _noalloc(L, torch, y2, p2)
    _check(icp_b200, got2, y2, p2)
    assert abs(got2[0] - 1.0) < 1e-12                      # noise-free rigid pair: s = 1
    # and call 1 again gives the bit-identical answer (fixed reduction order, scratch fully rewritten)
    assert np.array_equal(_noalloc(L, torch, y1, p1), got1)


def test_noalloc_rejects_what_it_cannot_serve():
    import torch
    import icp_b200
    L = icp_b200.load_library()
    torch.cuda.set_device(0)
    buf = torch.zeros(64, dtype=torch.float64, device="cuda")
    ptr = C.c_void_p(buf.data_ptr())
    rc = L.icp_find_alignment_noalloc(ptr, ptr, (1 << 22) + 1, ptr, None)          # beyond the stated maximum
    assert rc == icp_b200.ERR_INVALID and b"exceeds the maximum of 4194304" in L.icp_last_error()
    assert L.icp_find_alignment_noalloc(ptr, ptr, 0, ptr, None) == icp_b200.ERR_INVALID
    assert L.icp_find_alignment_noalloc(None, ptr, 10, ptr, None) == icp_b200.ERR_INVALID
