"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/icp_b200.h declares; without a GPU the product path fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "icp_b200.h")


def declared_symbols():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(icp_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for s in ["icp_closest", "icp_centroid", "icp_find_alignment", "icp_err_compute", "icp_run", "icp_set_model",
              "icp_set_scene", "icp_get_scene", "icp_closest_matrix_host", "icp_err_compute_host"]:
        assert s in syms


def test_library_exports_every_declared_symbol():
    import icp_b200
    L = icp_b200.load_library()
    for s in declared_symbols():
        assert hasattr(L, s), f"libicp_b200.so does not export {s}"
    assert sorted(icp_b200.EXPORTS) == declared_symbols()
    assert b"sm_100a" in L.icp_version()


def test_every_entry_point_cites_the_reference():
    txt = open(HEADER).read()
    for ref in ["src/cpu.cc:5-27", "src/cpu.cc:105-175", "src/cpu.cc:29-40", "src/cpu.cc:93-103", "src/cpu.cc:55-79",
                "src/GPU/gpu.hh:110", "src/cpu.hh:48-60"]:
        assert ref in txt


def test_no_cpu_fallback_without_device():
    import icp_b200
    L = icp_b200.load_library()
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = L.icp_create(C.byref(h), 0)
    assert rc == icp_b200.ERR_NO_DEVICE
    assert b"no CPU fallback" in L.icp_last_error()
    with pytest.raises(icp_b200.IcpError):
        icp_b200.closest_matrix([[0.0, 0, 0]], [[1.0, 0, 0]])
    with pytest.raises(icp_b200.IcpError) as e:            # the multi-device form fails the same way, and frees what it made
        icp_b200.Group([0, 1])
    assert e.value.code == icp_b200.ERR_NO_DEVICE
    for bad in ([], [0, 0], list(range(17))):
        with pytest.raises(icp_b200.IcpError) as e:
            icp_b200.Group(bad)
        assert e.value.code == icp_b200.ERR_INVALID


def test_product_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "iterative-closest-point_b200")
    for d, _, files in os.walk(pkg):
        if os.path.basename(d) in ("build", "lib", "bin", "__pycache__"):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".hh", ".h")) or f == "Makefile":
                src = open(os.path.join(d, f), errors="ignore").read()
                assert "liboracle" not in src and "import oracle" not in src and "from oracle" not in src, f
