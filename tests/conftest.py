import lzma
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "iterative-closest-point_b200")
GOLDEN = os.path.join(ROOT, "tests", "golden")
for p in (ROOT, PKG_DIR):
    if p not in sys.path:
        sys.path.insert(0, p)


# The library's default search mode is "auto" (brute force, the tree inside icp_run on small clouds).  The suite pins "brute" so
# that every test of the brute-force kernels keeps testing them on small inputs too; tree tests select the tree explicitly, and
# the auto rule has its own tests (which remove the variable).
os.environ.setdefault("ICP_NN", "brute")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def cloud_text_path(name, tmpdir):
    """Decompress tests/golden/clouds/<name>.txt.xz into tmpdir and return the .txt path."""
    dst = os.path.join(str(tmpdir), name + ".txt")
    if not os.path.exists(dst):
        with lzma.open(os.path.join(GOLDEN, "clouds", name + ".txt.xz"), "rb") as f:
            data = f.read()
        with open(dst, "wb") as g:
            g.write(data)
    return dst


@pytest.fixture(scope="session")
def cloud_dir(tmp_path_factory):
    return tmp_path_factory.mktemp("clouds")


@pytest.fixture(scope="session")
def clouds(cloud_dir):
    """name -> (n, 3) float64, parsed by the oracle's restatement of src/load.cc."""
    from oracle import oracle as O
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = O.load_matrix(cloud_text_path(name, cloud_dir))
        return cache[name]
    return get


@pytest.fixture(scope="session")
def golden():
    def get(case):
        return np.load(os.path.join(GOLDEN, case + ".npz"))
    return get


def synthetic_pair(n, seed=0, angle_deg=5.0, noise=1e-3, shuffle=True, shift=(0.05, -0.03, 0.02)):
    """Model ~ U[-1,1)^3; scene = R m + t + noise, optionally permuted: the SURVEY §8d recipe (splitmix64 / Box-Muller,
    icp_b200/synth.py — the same generator bench.py uses) with seeds (1000, 2000, 3000) + seed."""
    from icp_b200 import synth
    return synth.pair(n, 1000 + seed, 2000 + seed, 3000 + seed, angle_deg=angle_deg, noise=noise, shift=shift, shuffle=shuffle)
