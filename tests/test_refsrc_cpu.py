"""Source-level drop-in check: the reference's OWN translation units — src/main.cc (the `icp` driver),
src/GPU/main.cc (`icp-gpu`) and src/bench.cc (its google-benchmark harness) — compile UNMODIFIED against
host/compat/ (headers standing where src/cpu.hh, src/load.hh, src/GPU/gpu.hh, src/GPU/header.hh and
<benchmark/benchmark.h> stand) and link against libicp_b200.so.  The sources are compiled from where they lie under
/root/reference (nothing is copied), so the build half of this test only runs where that tree exists; the binaries it
leaves in bin/ travel to the GPU box, where tests/test_gpu_zz_refsrc.py runs them.  Without a GPU the binaries must
get as far as the reference does before its first computation and then fail loudly: no CPU fallback."""
import os
import subprocess

import pytest

from conftest import PKG_DIR, cloud_text_path

REF = "/root/reference"
BIN = os.path.join(PKG_DIR, "bin")
NAMES = ("ref_icp", "ref_icp_gpu", "ref_bench")
have_ref = os.path.exists(os.path.join(REF, "src", "bench.cc"))
have_bins = all(os.path.exists(os.path.join(BIN, n)) for n in NAMES)


def no_gpu_env():
    e = dict(os.environ)
    e["CUDA_VISIBLE_DEVICES"] = ""          # also on a GPU box: this file checks the no-device behaviour
    return e


@pytest.mark.skipif(not have_ref, reason="the reference tree is only present in the build container")
def test_reference_sources_compile_unmodified_and_link():
    r = subprocess.run(["make", "-C", PKG_DIR, "refsrc"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    for n in NAMES:
        path = os.path.join(BIN, n)
        assert os.path.exists(path)
        needed = subprocess.run(["readelf", "-d", path], capture_output=True, text=True).stdout
        assert "libicp_b200.so" in needed, n                 # the path is the library, not code compiled into the driver
    # the recipe feeds the reference's files themselves to the compiler (stdin), it does not copy or patch them
    mk = open(os.path.join(PKG_DIR, "Makefile")).read()
    for src in ("src/main.cc", "src/GPU/main.cc", "src/bench.cc"):
        assert f"< $(REF)/{src}" in mk
    assert not any(f.endswith((".cc", ".cu")) for _, _, fs in os.walk(os.path.join(PKG_DIR, "host", "compat")) for f in fs)


@pytest.mark.skipif(not have_bins, reason="bin/ref_* not built (needs the reference tree at build time)")
@pytest.mark.parametrize("name,usage", [("ref_icp", "./icp"), ("ref_icp_gpu", "./icp-gpu")])
def test_reference_drivers_keep_their_cli(tmp_path, cloud_dir, name, usage):
    exe = os.path.join(BIN, name)
    r = subprocess.run([exe], capture_output=True, text=True, cwd=tmp_path, timeout=60)
    assert r.returncode == 255                                # `return -1`, src/main.cc:11, src/GPU/main.cc:7
    assert r.stdout == f"Usage: {usage} [path_to_ref_cloud] [path_to_transform_cloud] [nb_iter]\n"
    r = subprocess.run([exe, "nope.txt", "nope2.txt", "3"], capture_output=True, text=True, cwd=tmp_path, timeout=60)
    assert r.returncode == 2 and "could not be opened" in r.stderr          # src/load.cc:11-14
    ref, scene = cloud_text_path("cow_ref", cloud_dir), cloud_text_path("cow_tr1", cloud_dir)
    r = subprocess.run([exe, ref, scene, "3"], capture_output=True, text=True, cwd=tmp_path, timeout=120, env=no_gpu_env())
    assert r.returncode == 3                                  # icpb::die: loud, not a silent CPU path
    assert "no usable CUDA device (this library has no CPU fallback)" in r.stderr
    assert not (tmp_path / "output.txt").exists()


@pytest.mark.skipif(not have_bins, reason="bin/ref_* not built (needs the reference tree at build time)")
def test_reference_bench_harness_fails_loudly_without_a_device(tmp_path, cloud_dir):
    d = tmp_path / "data_students"                            # REF_PATH / SCENE_PATH of src/bench.cc:7-8 are relative
    d.mkdir()
    for n in ("cow_ref", "cow_tr1"):
        os.rename(cloud_text_path(n, tmp_path), d / (n + ".txt"))
    r = subprocess.run([os.path.join(BIN, "ref_bench")], capture_output=True, text=True, cwd=tmp_path, timeout=120,
                       env=no_gpu_env())
    assert r.returncode == 3 and "no CPU fallback" in r.stderr
    assert "[load] opening data_students/cow_ref.txt" in r.stderr
