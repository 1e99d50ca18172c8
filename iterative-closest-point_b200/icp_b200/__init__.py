"""icp_b200 — Python host mirror of the reference's ICP interface over the C-ABI (include/icp_b200.h).

This module holds no arithmetic: every operator is a call into lib/libicp_b200.so (hand-written CUDA
for sm_100a).  There is no CPU fallback — if the library or a CUDA device is missing the calls raise.

Reference interface mirrored (file:line in yassram/iterative-closest-point):
  ICP(m, p, max_iter)                 CPU::ICP ctor                 src/cpu.hh:48-60
  ICP.find_corresponding()            src/cpu.cc:55-79 / GPU::ICP::find_corresponding_opti src/GPU/gpu.cc:52-83
  ICP.find_alignment(Y)               src/cpu.cc:105-175
  ICP.new_p                           src/cpu.hh:97
  closest_matrix(p, m)                CPU::closest_matrix           src/cpu.cc:5-27
  err_compute / err_compute_alignment src/cpu.cc:29-40, 93-103
  compute_Y_w_opti, compute_err_w, substract_col_w, y_p_norm_w, compute_distance_w_naive
                                      src/GPU/gpu.hh:110-116
Clouds are (n, 3) float64 arrays — the memory image of the reference's column-major 3 x n MatrixXd.
"""
import ctypes as C
import os
import sys

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_PKG), "lib", "libicp_b200.so")

ICP_OK = 0
ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_SIZE_MISMATCH, ERR_TOO_FEW, ERR_COMM = -1, -2, -3, -4, -5, -6

EXPORTS = [
    "icp_create", "icp_destroy", "icp_last_error", "icp_version", "icp_set_model", "icp_set_scene",
    "icp_get_scene", "icp_reset_scene", "icp_closest", "icp_centroid", "icp_find_alignment", "icp_err_compute",
    "icp_get_transform", "icp_run", "icp_get_indices", "icp_get_stats", "icp_closest_matrix_host",
    "icp_nearest_index_host", "icp_find_alignment_host", "icp_err_compute_host", "icp_subtract_col_host", "icp_y_p_norm_host",
    "icp_comm_unique_id", "icp_comm_init", "icp_comm_world", "icp_comm_peer_handle", "icp_comm_peer_attach", "icp_get_trace", "icp_get_block_trace", "icp_batch_run", "icp_batch_kernels_launched", "icp_release_cached",
    "icp_set_nn_mode", "icp_get_nn_mode",
    "icp_group_create", "icp_group_destroy", "icp_group_size", "icp_group_member", "icp_group_set_model", "icp_group_set_scene",
    "icp_group_reset_scene", "icp_group_set_nn_mode", "icp_group_closest", "icp_group_centroid", "icp_group_find_alignment",
    "icp_group_err_compute", "icp_group_run", "icp_group_get_scene", "icp_group_get_indices", "icp_group_get_transform",
    "icp_group_get_stats", "icp_batch_run_multi", "icp_find_alignment_noalloc",
]


class IcpStats(C.Structure):
    _fields_ = [("np", C.c_int64), ("nm", C.c_int64), ("nn_calls", C.c_int64), ("nn_pairs", C.c_int64),
                ("rerank_full", C.c_int64), ("rerank_chunk", C.c_int64), ("n_splits", C.c_int32),
                ("tile_points", C.c_int32), ("chunk_points", C.c_int32), ("kernels_launched", C.c_int32),
                ("ms_filter", C.c_float), ("ms_total", C.c_float), ("ms_refine", C.c_float), ("ms_exact", C.c_float),
                ("ms_align", C.c_float), ("reserved_", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class IcpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"icp_b200 error {code}: {msg}")
        self.code = code


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_lib = None


def load_library(path=None):
    """dlopen libicp_b200.so and declare the prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = path or os.environ.get("ICP_B200_LIB") or LIB_PATH      # ICP_B200_LIB: A/B runs of two builds
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: build it with `make -C {os.path.dirname(_PKG)}` "
                          "(there is no CPU fallback)")
    L = C.CDLL(path)
    vp = C.c_void_p
    L.icp_last_error.restype = C.c_char_p
    L.icp_version.restype = C.c_char_p
    L.icp_create.argtypes = [C.POINTER(vp), C.c_int]
    L.icp_destroy.argtypes = [vp]
    L.icp_destroy.restype = None
    L.icp_set_model.argtypes = [vp, _dp, C.c_int64]
    L.icp_set_scene.argtypes = [vp, _dp, C.c_int64]
    L.icp_get_scene.argtypes = [vp, _dp]
    L.icp_reset_scene.argtypes = [vp]
    L.icp_set_nn_mode.argtypes = [vp, C.c_int]
    L.icp_get_nn_mode.argtypes = [vp]
    L.icp_closest.argtypes = [vp, _ip, _dp]
    L.icp_centroid.argtypes = [vp, _dp, _dp]
    L.icp_find_alignment.argtypes = [vp, _dp, _dp, _dp, _dp]
    L.icp_err_compute.argtypes = [vp, C.c_int, _dp]
    L.icp_get_transform.argtypes = [vp, _dp, _dp, _dp]
    L.icp_run.argtypes = [vp, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int), _dp]
    L.icp_get_indices.argtypes = [vp, _ip]
    L.icp_get_stats.argtypes = [vp, C.POINTER(IcpStats)]
    L.icp_closest_matrix_host.argtypes = [_dp, C.c_int64, _dp, C.c_int64, _ip, _dp]
    L.icp_nearest_index_host.argtypes = [_dp, C.c_int64, _dp, _ip]
    L.icp_find_alignment_host.argtypes = [_dp, _dp, C.c_int64, _dp, _dp, _dp, _dp]
    L.icp_err_compute_host.argtypes = [_dp, _dp, C.c_int64, C.c_int, _dp, _dp, _dp]
    L.icp_subtract_col_host.argtypes = [_dp, C.c_int64, _dp, _dp]
    L.icp_y_p_norm_host.argtypes = [_dp, _dp, C.c_int64, _dp, _dp]
    L.icp_comm_unique_id.argtypes = [C.c_char_p, C.c_void_p]
    L.icp_comm_init.argtypes = [vp, C.c_char_p, C.c_void_p, C.c_int, C.c_int]
    L.icp_comm_world.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.icp_get_trace.argtypes = [vp, _dp]
    L.icp_get_block_trace.argtypes = [vp, _dp]
    L.icp_batch_run.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_void_p, C.POINTER(C.c_int64), C.c_void_p,
                                C.POINTER(C.c_int64), C.c_int, C.c_double, C.c_int, _ip, _dp, C.c_void_p, _dp]
    L.icp_find_alignment_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    L.icp_batch_kernels_launched.restype = C.c_longlong
    L.icp_batch_run_multi.argtypes = [C.POINTER(C.c_int), C.c_int] + list(L.icp_batch_run.argtypes[1:])
    L.icp_group_create.argtypes = [C.POINTER(vp), C.POINTER(C.c_int), C.c_int]
    L.icp_group_destroy.argtypes = [vp]
    L.icp_group_destroy.restype = None
    L.icp_group_size.argtypes = [vp]
    L.icp_group_member.argtypes = [vp, C.c_int]
    L.icp_group_member.restype = vp
    L.icp_group_set_model.argtypes = [vp, _dp, C.c_int64]
    L.icp_group_set_scene.argtypes = [vp, _dp, C.c_int64]
    L.icp_group_reset_scene.argtypes = [vp]
    L.icp_group_set_nn_mode.argtypes = [vp, C.c_int]
    L.icp_group_closest.argtypes = [vp, _ip, _dp]
    L.icp_group_centroid.argtypes = [vp, _dp, _dp]
    L.icp_group_find_alignment.argtypes = [vp, _dp, _dp, _dp, _dp]
    L.icp_group_err_compute.argtypes = [vp, C.c_int, _dp]
    L.icp_group_run.argtypes = [vp, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int), _dp]
    L.icp_group_get_scene.argtypes = [vp, _dp]
    L.icp_group_get_indices.argtypes = [vp, _ip]
    L.icp_group_get_transform.argtypes = [vp, _dp, _dp, _dp]
    L.icp_group_get_stats.argtypes = [vp, C.POINTER(IcpStats)]
    L.icp_comm_peer_handle.argtypes = [vp, C.c_void_p]
    L.icp_comm_peer_attach.argtypes = [vp, C.c_void_p, C.c_int, C.c_int]
    _lib = L
    return L


def _ck(rc):
    if rc != ICP_OK:
        raise IcpError(rc, load_library().icp_last_error().decode())


def _pts(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if a.ndim != 2 or a.shape[1] != 3:
        raise ValueError("clouds are (n, 3) float64 arrays (the reference's 3 x n column-major MatrixXd)")
    return a


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def torch_nccl_path():
    """Path of the libnccl.so.2 that torch.distributed uses, so both share one NCCL."""
    try:
        import nvidia.nccl  # type: ignore
        d = os.path.join(list(nvidia.nccl.__path__)[0], "lib", "libnccl.so.2")
        if os.path.exists(d):
            return d
    except Exception:
        pass
    return None


class Context:
    """Thin owner of an icp_ctx (device-resident model, scene, scratch)."""

    def __init__(self, device=0):
        self._L = load_library()
        self._h = C.c_void_p()
        _ck(self._L.icp_create(C.byref(self._h), int(device)))
        self.np = 0
        self.nm = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.icp_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_model(self, m):
        m = _pts(m)
        _ck(self._L.icp_set_model(self._h, _d(m), len(m)))
        self.nm = len(m)

    def set_scene(self, p):
        p = _pts(p)
        _ck(self._L.icp_set_scene(self._h, _d(p), len(p)))
        self.np = len(p)

    def reset_scene(self):
        _ck(self._L.icp_reset_scene(self._h))

    def set_nn_mode(self, mode):
        """'auto' (default: brute force, the tree inside run() on small clouds), 'brute' (north_star) or 'tree' (exact
        bounding-box hierarchy); same results.  nn_mode() reports what the last search used."""
        _ck(self._L.icp_set_nn_mode(self._h, {"brute": 0, "tree": 1, "auto": 2}[mode]))

    def nn_mode(self):
        return ("brute", "tree")[self._L.icp_get_nn_mode(self._h)]

    def get_scene(self, out=None):
        """new_p (src/cpu.hh:97) as (np, 3) float64; `out`: a C-contiguous float64 buffer of that shape to fill (e.g. pinned)."""
        if out is None:
            out = np.empty((self.np, 3))
        assert out.dtype == np.float64 and out.shape == (self.np, 3) and out.flags.c_contiguous
        _ck(self._L.icp_get_scene(self._h, _d(out)))
        return out

    def closest(self, want_idx=True, want_Y=True):
        idx = np.empty(self.np, np.int32) if want_idx else None
        Y = np.empty((self.np, 3)) if want_Y else None
        _ck(self._L.icp_closest(self._h, _i(idx) if want_idx else None, _d(Y) if want_Y else None))
        return idx, Y

    def indices(self):
        idx = np.empty(self.np, np.int32)
        _ck(self._L.icp_get_indices(self._h, _i(idx)))
        return idx

    def centroid(self):
        a, b = np.empty(3), np.empty(3)
        _ck(self._L.icp_centroid(self._h, _d(a), _d(b)))
        return a, b

    def find_alignment(self):
        s = C.c_double(); e = C.c_double(); R = np.empty(9); t = np.empty(3)
        _ck(self._L.icp_find_alignment(self._h, C.byref(s), _d(R), _d(t), C.byref(e)))
        return dict(s=s.value, R=R.reshape(3, 3), t=t, err=e.value)

    def transform(self):
        s = C.c_double(); R = np.empty(9); t = np.empty(3)
        _ck(self._L.icp_get_transform(self._h, C.byref(s), _d(R), _d(t)))
        return s.value, R.reshape(3, 3), t

    def err_compute(self, in_place=True):
        e = C.c_double()
        _ck(self._L.icp_err_compute(self._h, int(bool(in_place)), C.byref(e)))
        return e.value

    def run(self, max_iter, threshold=1e-5, check_sizes=True):
        it = C.c_int()
        hist = np.zeros(max(int(max_iter), 1))
        _ck(self._L.icp_run(self._h, int(max_iter), float(threshold), int(bool(check_sizes)), C.byref(it), _d(hist)))
        return it.value, hist[:it.value].copy()

    def stats(self):
        st = IcpStats()
        _ck(self._L.icp_get_stats(self._h, C.byref(st)))
        return st.as_dict()

    def trace(self):
        t = np.zeros(10)
        _ck(self._L.icp_get_trace(self._h, _d(t)))
        return t

    def block_trace(self):
        """(32, 5): SM id, entry, barrier passed, solve done, exit (us after block 0's entry) of the first 32 blocks."""
        t = np.zeros(160)
        _ck(self._L.icp_get_block_trace(self._h, _d(t)))
        return t.reshape(32, 5)

    def comm_init(self, id128, rank, world, nccl_path=None):
        path = nccl_path or torch_nccl_path()
        buf = C.create_string_buffer(bytes(id128), 128)
        _ck(self._L.icp_comm_init(self._h, path.encode() if path else None, buf, int(rank), int(world)))


    def comm_peer_handle(self):
        """64-byte CUDA IPC handle of this rank's mailbox (all-gather these, then comm_peer_attach)."""
        buf = C.create_string_buffer(64)
        _ck(self._L.icp_comm_peer_handle(self._h, buf))
        return buf.raw

    def comm_peer_attach(self, handles, rank, world):
        """handles: list of `world` 64-byte handles in rank order."""
        blob = b"".join(handles)
        assert len(blob) == 64 * world
        buf = C.create_string_buffer(blob, len(blob))
        _ck(self._L.icp_comm_peer_attach(self._h, buf, int(rank), int(world)))


class Group:
    """One process, several GPUs (icp_group_*): the model replicated, the scene cut into contiguous blocks, the
    per-iteration exchange inside the alignment kernels through peer memory.  Mirrors Context call for call."""

    def __init__(self, devices):
        self._L = load_library()
        self._h = C.c_void_p()
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        _ck(self._L.icp_group_create(C.byref(self._h), devs, len(devices)))
        self.devices = list(devices)
        self.np = 0
        self.nm = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.icp_group_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_model(self, m):
        m = _pts(m)
        _ck(self._L.icp_group_set_model(self._h, _d(m), len(m)))
        self.nm = len(m)

    def set_scene(self, p):
        p = _pts(p)
        _ck(self._L.icp_group_set_scene(self._h, _d(p), len(p)))
        self.np = len(p)

    def reset_scene(self):
        _ck(self._L.icp_group_reset_scene(self._h))

    def set_nn_mode(self, mode):
        _ck(self._L.icp_group_set_nn_mode(self._h, {"brute": 0, "tree": 1, "auto": 2}[mode]))

    def closest(self, want_idx=True, want_Y=True):
        idx = np.empty(self.np, np.int32) if want_idx else None
        Y = np.empty((self.np, 3)) if want_Y else None
        _ck(self._L.icp_group_closest(self._h, _i(idx) if want_idx else None, _d(Y) if want_Y else None))
        return idx, Y

    def run(self, max_iter, threshold=1e-5, check_sizes=True):
        it = C.c_int()
        hist = np.zeros(max(int(max_iter), 1))
        _ck(self._L.icp_group_run(self._h, int(max_iter), float(threshold), int(bool(check_sizes)), C.byref(it), _d(hist)))
        return it.value, hist[:it.value].copy()

    def centroid(self):
        a, b = np.empty(3), np.empty(3)
        _ck(self._L.icp_group_centroid(self._h, _d(a), _d(b)))
        return a, b

    def find_alignment(self):
        s = np.zeros(1); e = np.zeros(1); R = np.empty(9); t = np.empty(3)
        _ck(self._L.icp_group_find_alignment(self._h, _d(s), _d(R), _d(t), _d(e)))
        return dict(s=float(s[0]), R=R.reshape(3, 3), t=t, err=float(e[0]))

    def err_compute(self, in_place=True):
        e = np.zeros(1)
        _ck(self._L.icp_group_err_compute(self._h, int(bool(in_place)), _d(e)))
        return float(e[0])

    def get_scene(self, out=None):
        if out is None:
            out = np.empty((self.np, 3))
        assert out.dtype == np.float64 and out.shape == (self.np, 3) and out.flags.c_contiguous
        _ck(self._L.icp_group_get_scene(self._h, _d(out)))
        return out

    def indices(self):
        idx = np.empty(self.np, np.int32)
        _ck(self._L.icp_group_get_indices(self._h, _i(idx)))
        return idx

    def transform(self):
        s = np.zeros(1); R = np.empty(9); t = np.empty(3)
        _ck(self._L.icp_group_get_transform(self._h, _d(s), _d(R), _d(t)))
        return float(s[0]), R.reshape(3, 3), t

    def stats(self):
        st = IcpStats()
        _ck(self._L.icp_group_get_stats(self._h, C.byref(st)))
        return st.as_dict()

    def member_stats(self, i):
        st = IcpStats()
        _ck(self._L.icp_get_stats(self._L.icp_group_member(self._h, int(i)), C.byref(st)))
        return st.as_dict()


def batch_run(models, scenes, max_iter=30, threshold=1e-5, check_sizes=True, device=0, lanes=3, want_scene=True, devices=None):
    """Registers scenes[k] onto models[k] for every k (independent pairs, reference loop src/cpu.cc:55-79).
    devices: a list of GPUs of this process to deal the pairs to (icp_batch_run_multi); default: `device` alone.
    Returns dict(iters, err, srt, kernels_launched[, new_p])."""
    L = load_library()
    n = len(models)
    ms = [_pts(a) for a in models]
    ps = [_pts(a) for a in scenes]
    PP = C.POINTER(C.c_double) * max(n, 1)
    mp = PP(*[_d(a) for a in ms])
    pp = PP(*[_d(a) for a in ps])
    nm = (C.c_int64 * max(n, 1))(*[len(a) for a in ms])
    npts = (C.c_int64 * max(n, 1))(*[len(a) for a in ps])
    iters = np.zeros(max(n, 1), dtype=np.int32)
    err = np.zeros(max(n, 1))
    srt = np.zeros((max(n, 1), 13))
    outs = [np.empty_like(a) for a in ps] if want_scene else []
    op = PP(*[_d(a) for a in outs]) if want_scene else None
    if devices is not None and len(devices) > 1:
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        _ck(L.icp_batch_run_multi(devs, len(devices), int(lanes), n, mp, nm, pp, npts, int(max_iter), float(threshold),
                                  int(bool(check_sizes)), _i(iters), _d(err), op, _d(srt)))
    else:
        _ck(L.icp_batch_run(int(devices[0] if devices else device), int(lanes), n, mp, nm, pp, npts, int(max_iter), float(threshold),
                            int(bool(check_sizes)), _i(iters), _d(err), op, _d(srt)))
    res = dict(iters=iters[:n], err=err[:n], srt=srt[:n], kernels_launched=int(L.icp_batch_kernels_launched()))
    if want_scene:
        res["new_p"] = outs
    return res


def shard_range(n, rank, world):
    """Contiguous block of scene points held by `rank` (scene sharded, model replicated)."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


# Layout of the per-iteration exchange (csrc/icp_common.cuh MOM_N): all-reduced with SUM.
MOMENT_FIELDS = ("n", "sum_a[3]", "sum_b[3]", "sum_abT[9]", "sum_aa", "sum_bb")   # a = p - c0, b = y - c0
MOMENT_DOUBLES = 18


def comm_unique_id(nccl_path=None):
    path = nccl_path or torch_nccl_path()
    buf = C.create_string_buffer(128)
    _ck(load_library().icp_comm_unique_id(path.encode() if path else None, buf))
    return buf.raw


# ---- reference-shaped free functions (src/cpu.cc, src/GPU/gpu.hh:110-116) -----------------------------

def closest_matrix(p, m, return_index=False):
    """CPU::closest_matrix / compute_Y_w_opti: Y[j] = m[argmin_k |p_j - m_k|], first minimum."""
    p, m = _pts(p), _pts(m)
    idx = np.empty(len(p), np.int32)
    Y = np.empty_like(p)
    _ck(load_library().icp_closest_matrix_host(_d(m), len(m), _d(p), len(p), _i(idx), _d(Y)))
    return (Y, idx) if return_index else Y


compute_Y_w_opti = closest_matrix


def compute_distance_w_naive(m, pi):
    """src/GPU/compute.cu:279-308: index of the model point nearest to the single point pi."""
    m = _pts(m)
    pi = np.ascontiguousarray(pi, np.float64).reshape(3)
    k = C.c_int32()
    _ck(load_library().icp_nearest_index_host(_d(m), len(m), _d(pi), C.byref(k)))
    return k.value


def compute_err_w(Y, p, in_place, sr, t):
    """src/GPU/compute.cu:348-379 -> (err, p) ; p is the transformed cloud when in_place."""
    Y, p = _pts(Y), _pts(p).copy()
    sr = np.ascontiguousarray(sr, np.float64).reshape(9)
    t = np.ascontiguousarray(t, np.float64).reshape(3)
    e = C.c_double()
    _ck(load_library().icp_err_compute_host(_d(Y), _d(p), len(p), int(bool(in_place)), _d(sr), _d(t), C.byref(e)))
    return e.value, p


def find_alignment(p, y):
    """ICP::find_alignment(Y) on explicit pairs (src/cpu.cc:105-175) -> dict(s, R, t, err)."""
    p, y = _pts(p), _pts(y)
    s = C.c_double(); e = C.c_double(); R = np.empty(9); t = np.empty(3)
    _ck(load_library().icp_find_alignment_host(_d(y), _d(p), len(p), C.byref(s), _d(R), _d(t), C.byref(e)))
    return dict(s=s.value, R=R.reshape(3, 3), t=t, err=e.value)


def err_compute(p, s, R, t, Y):
    """CPU::err_compute (src/cpu.cc:29-40) -> (err, new_p)."""
    return compute_err_w(Y, p, True, float(s) * np.asarray(R, np.float64), t)


def err_compute_alignment(p, s, R, t, y):
    """CPU::err_compute_alignment (src/cpu.cc:93-103) -> err."""
    return compute_err_w(y, p, False, float(s) * np.asarray(R, np.float64), t)[0]


def substract_col_w(M, m):
    M = _pts(M)
    m = np.ascontiguousarray(m, np.float64).reshape(3)
    out = np.empty_like(M)
    _ck(load_library().icp_subtract_col_host(_d(M), len(M), _d(m), _d(out)))
    return out


def y_p_norm_w(y, p):
    y, p = _pts(y), _pts(p)
    d_caps, sp = C.c_double(), C.c_double()
    _ck(load_library().icp_y_p_norm_host(_d(y), _d(p), len(p), C.byref(d_caps), C.byref(sp)))
    return d_caps.value, sp.value


class ICP:
    """Mirror of CPU::ICP / GPU::ICP (src/cpu.hh:45-114): same ctor, find_corresponding(),
    find_alignment(Y), public new_p; error behaviour of alignement_check surfaces as SystemExit(-1)
    with the reference's messages on stderr (src/cpu.cc:42-53)."""

    threshold = 1e-5   # src/cpu.hh:113

    def __init__(self, m, p, max_iter, device=0, log=True):
        self.m, self.p = _pts(m), _pts(p)
        self.max_iter = int(max_iter)
        self.log = log
        self.ctx = Context(device)
        self.ctx.set_model(self.m)
        self.ctx.set_scene(self.p)
        self.new_p = self.p.copy()
        self.s, self.r, self.t = 1.0, np.eye(3), np.zeros(3)
        self.iterations, self.errors = 0, np.zeros(0)

    def find_corresponding(self, check_sizes=True):
        try:
            it, hist = self.ctx.run(self.max_iter, self.threshold, check_sizes)
        except IcpError as e:
            if e.code == ERR_SIZE_MISMATCH:
                sys.stderr.write("[error] Point sets need to have the same number of points.\n")
                raise SystemExit(-1)
            if e.code == ERR_TOO_FEW:
                sys.stderr.write("[error] Need at least 4 point pairs\n")
                raise SystemExit(-1)
            raise
        if self.log:
            for i, e in enumerate(hist):
                sys.stderr.write("[ICP] iteration number %d | error value = %s\n" % (i, format(e, ".6g")))
        self.iterations, self.errors = it, hist
        self.new_p = self.ctx.get_scene()
        self.s, self.r, self.t = self.ctx.transform()
        return it

    find_corresponding_opti = find_corresponding

    def closest_matrix(self):
        return self.ctx.closest()[1]

    def find_alignment(self, Y=None):
        """Uses the correspondences of the last closest pass (Y is accepted for signature parity)."""
        r = self.ctx.find_alignment()
        self.s, self.r, self.t = r["s"], r["R"], r["t"]
        return r["err"]
