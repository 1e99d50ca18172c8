"""GPU test of the `icp [ref] [scene] [nb_max_iter]` driver (reference src/main.cc:5-25): same stdout /
stderr text, same output.txt bytes as the oracle's restatement of the reference's loader, loop and writer."""
import os
import subprocess

import numpy as np
import pytest

from conftest import PKG_DIR, cloud_text_path
from oracle import oracle as O

pytestmark = pytest.mark.gpu
ICP_BIN = os.path.join(PKG_DIR, "bin", "icp")


def run_icp(args, cwd, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([ICP_BIN] + args, cwd=str(cwd), capture_output=True, text=True, timeout=300, env=e)


def test_usage_and_missing_file(tmp_path):
    r = run_icp([], tmp_path)
    assert r.returncode == 255 and r.stdout == "Usage: ./icp [path_to_ref_cloud] [path_to_transform_cloud] [nb_iter]\n"
    r = run_icp(["nope.txt", "nope2.txt", "3"], tmp_path)
    assert r.returncode == 2                                    # src/load.cc:11-14
    assert "[load] opening nope.txt" in r.stderr and "could not be opened" in r.stderr


def test_cow_end_to_end(tmp_path, cloud_dir):
    ref, scene = cloud_text_path("cow_ref", cloud_dir), cloud_text_path("cow_tr1", cloud_dir)
    r = run_icp([ref, scene, "10"], tmp_path)
    assert r.returncode == 0, r.stderr
    lines = r.stderr.splitlines()
    assert lines[0] == f"[load] opening {ref}" and lines[1] == "[load] loading file into matrix"
    it_lines = [l for l in lines if l.startswith("[ICP] iteration number")]
    ro = O.icp_run(O.load_matrix(ref), O.load_matrix(scene), 10)
    assert len(it_lines) == ro["iters"] == 7                     # the reference's pinned known answer
    for i, l in enumerate(it_lines):
        assert l.startswith(f"[ICP] iteration number {i} | error value = ")
        assert abs(float(l.split("=")[1]) - ro["err"][i]) <= 2e-5 * ro["err"][i]   # 6 significant digits printed
    assert lines[-1] == '[output] output file "output.txt" was generated.'
    O.write_matrix(str(tmp_path / "oracle_output.txt"), ro["new_p"])
    got = (tmp_path / "output.txt").read_text().splitlines()
    want = (tmp_path / "oracle_output.txt").read_text().splitlines()
    assert got[0] == want[0] == "Points_0,Points_1,Points_2" and len(got) == len(want) == 2904
    g = np.array([[float(x) for x in l.split(",")] for l in got[1:]])
    w = np.array([[float(x) for x in l.split(",")] for l in want[1:]])
    assert np.abs(g - w).max() <= 1e-5 * np.abs(w).max()
    assert sum(a != b for a, b in zip(got, want)) <= 5             # 6-digit text: identical but for rounding flips


def test_size_mismatch_exits_like_reference(tmp_path, cloud_dir):
    ref, scene = cloud_text_path("bun000", cloud_dir), cloud_text_path("bun045", cloud_dir)
    r = run_icp([ref, scene, "2"], tmp_path)
    assert r.returncode == 255                                   # exit(-1), src/cpu.cc:44-47
    assert "[error] Point sets need to have the same number of points." in r.stderr
    r = run_icp([ref, scene, "2"], tmp_path, env={"ICP_ALLOW_UNEQUAL": "1"})
    assert r.returncode == 0 and r.stderr.count("[ICP] iteration number") == 2


def test_tree_search_driver_writes_the_same_bytes(tmp_path, cloud_dir):
    # ICP_NN=tree (read by icp_create) changes how the correspondences are found, not what they are: the driver's
    # stderr lines and output.txt must be byte-identical to the brute-force run
    ref, scene = cloud_text_path("cow_ref", cloud_dir), cloud_text_path("cow_tr1", cloud_dir)
    (tmp_path / "a").mkdir()
    (tmp_path / "b").mkdir()
    ra = run_icp([ref, scene, "10"], tmp_path / "a", env={"ICP_NN": "brute"})
    rb = run_icp([ref, scene, "10"], tmp_path / "b", env={"ICP_NN": "tree"})
    assert ra.returncode == rb.returncode == 0
    assert ra.stderr == rb.stderr and ra.stdout == rb.stdout
    assert (tmp_path / "a" / "output.txt").read_bytes() == (tmp_path / "b" / "output.txt").read_bytes()
    (tmp_path / "c").mkdir()                           # the library's default ("auto": the tree on a cloud of cow's size)
    rc = run_icp([ref, scene, "10"], tmp_path / "c", env={"ICP_NN": "auto"})
    assert rc.returncode == 0 and rc.stderr == ra.stderr
    assert (tmp_path / "a" / "output.txt").read_bytes() == (tmp_path / "c" / "output.txt").read_bytes()


def test_two_gpu_driver_writes_the_same_bytes(tmp_path, cloud_dir):
    # ICP_DEVICES=0,1: the drop-in `icp` driver shards the scene over two GPUs of its own process (icp_group_*); the
    # 18 moment sums are then added in a different order (per-device partials, rank order), so the printed 6-digit
    # error values and output.txt must agree with the one-GPU run (byte for byte unless a 6th digit sits on a rounding edge)
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ref, scene = cloud_text_path("cow_ref", cloud_dir), cloud_text_path("cow_tr1", cloud_dir)
    (tmp_path / "a").mkdir()
    (tmp_path / "b").mkdir()
    ra = run_icp([ref, scene, "10"], tmp_path / "a")
    rb = run_icp([ref, scene, "10"], tmp_path / "b", env={"ICP_DEVICES": "0,1"})
    assert ra.returncode == rb.returncode == 0, rb.stderr
    la, lb = ra.stderr.splitlines(), rb.stderr.splitlines()
    assert len(la) == len(lb) and sum(x != y for x, y in zip(la, lb)) <= 1
    a = (tmp_path / "a" / "output.txt").read_text().splitlines()
    b = (tmp_path / "b" / "output.txt").read_text().splitlines()
    assert len(a) == len(b) == 2904 and sum(x != y for x, y in zip(a, b)) <= 5
