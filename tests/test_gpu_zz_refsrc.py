"""GPU run of the reference's own drivers and benchmark harness built against the B200 path (bin/ref_icp = src/main.cc,
bin/ref_icp_gpu = src/GPU/main.cc, bin/ref_bench = src/bench.cc; see tests/test_refsrc_cpu.py for how they are
built).  They must behave like host/main.cc's `icp`, which tests/test_gpu_driver.py checks against the oracle."""
import os
import re
import subprocess

import pytest

from conftest import PKG_DIR, cloud_text_path

pytestmark = pytest.mark.gpu
BIN = os.path.join(PKG_DIR, "bin")
have_bins = all(os.path.exists(os.path.join(BIN, n)) for n in ("ref_icp", "ref_icp_gpu", "ref_bench", "icp"))
needs_bins = pytest.mark.skipif(not have_bins, reason="bin/ref_* not built (needs the reference tree at build time)")

# the 14 captures of src/bench.cc:391-445, in registration order
CAPTURES = [
    "BM_CPU_closest_matrix/cpu_closest_matrix", "BM_GPU_closest_matrix_naive/naive_gpu_closest_matrix",
    "BM_GPU_closest_matrix_opti/opti_gpu_closest_matrix", "BM_CPU_find_alignment/cpu_find_alignment",
    "BM_GPU_find_alignment/gpu_find_alignment", "BM_CPU_Compute_centroid/cpu_compute_centroid",
    "BM_GPU_Compute_centroid/gpu_compute_centroid", "BM_CPU_Err_compute/cpu_err_compute",
    "BM_GPU_Err_compute/gpu_err_compute", "BM_CPU_Err_compute_alignment/cpu_err_compute_alignment",
    "BM_GPU_Err_compute/gpu_err_compute_alignment", "BM_CPU_Find_corresponding/cpu_loop",
    "BM_GPU_Find_corresponding_naive/naive_gpu_loop", "BM_GPU_Find_corresponding_opti/opti_gpu_loop"]


def run(exe, args, cwd, env=None, timeout=300):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([os.path.join(BIN, exe)] + args, cwd=str(cwd), capture_output=True, text=True, timeout=timeout, env=e)


@needs_bins
@pytest.mark.parametrize("exe", ["ref_icp", "ref_icp_gpu"])
def test_reference_main_gives_the_same_run_as_the_driver(tmp_path, cloud_dir, exe):
    ref, scene = cloud_text_path("cow_ref", cloud_dir), cloud_text_path("cow_tr1", cloud_dir)
    (tmp_path / "a").mkdir()
    (tmp_path / "b").mkdir()
    ra = run("icp", [ref, scene, "10"], tmp_path / "a")
    rb = run(exe, [ref, scene, "10"], tmp_path / "b")
    assert ra.returncode == 0 and rb.returncode == 0, rb.stderr
    assert rb.stderr.count("[ICP] iteration number") == 7          # the reference's pinned known answer
    assert ra.stderr == rb.stderr and ra.stdout == rb.stdout
    assert (tmp_path / "a" / "output.txt").read_bytes() == (tmp_path / "b" / "output.txt").read_bytes()


@needs_bins
def test_reference_main_refuses_unequal_clouds(tmp_path, cloud_dir):
    ref, scene = cloud_text_path("bun000", cloud_dir), cloud_text_path("bun045", cloud_dir)
    r = run("ref_icp", [ref, scene, "2"], tmp_path)
    assert r.returncode == 255                                     # exit(-1), src/cpu.cc:44-47
    assert "[error] Point sets need to have the same number of points." in r.stderr


@needs_bins
def test_reference_bench_harness_runs_all_its_captures(tmp_path):
    d = tmp_path / "data_students"                                 # REF_PATH / SCENE_PATH, src/bench.cc:7-8
    d.mkdir()
    for n in ("cow_ref", "cow_tr1"):
        os.rename(cloud_text_path(n, tmp_path), d / (n + ".txt"))
    r = run("ref_bench", [], tmp_path, env={"ICP_BENCH_MIN_TIME": "0.02"}, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    rows = [l for l in r.stdout.splitlines() if l.startswith("BM_")]
    assert [l.split("/real_time")[0] for l in rows] == CAPTURES
    for l in rows:
        m = re.search(r"/real_time\s+([0-9.]+) ms\s+(\d+) frame_rate=", l)
        assert m and float(m.group(1)) >= 0.0 and int(m.group(2)) >= 1, l
    # the loops are the reference's: max_iter 20 on cow_ref / cow_tr1 stops after 7 iterations every time
    assert r.stderr.count("[ICP] iteration number 6 |") > 0 and "[ICP] iteration number 7 |" not in r.stderr
