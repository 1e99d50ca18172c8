"""The synthetic clouds of SURVEY §8d (configs 4 and 5) are defined by a recipe, not by a numpy RNG version:
known answers of splitmix64, bit-exact model coordinates, and the statistical shape of the scene."""
import zlib

import numpy as np

from icp_b200 import synth


def test_splitmix64_known_answers():
    # published test vectors of splitmix64 (seed 0 and 1234567)
    assert [int(x) for x in synth.splitmix64(0, 3)] == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F]
    assert [int(x) for x in synth.splitmix64(1234567, 2)] == [6457827717110365317, 3203168211198807973]


def test_config4_model_is_bit_reproducible():
    m, p = synth.config4(1 << 12)
    assert m.shape == p.shape == (4096, 3) and m.flags.c_contiguous and p.flags.c_contiguous
    assert zlib.crc32(m.tobytes()) == 37481346                  # pure integer -> double arithmetic: exact everywhere
    u = int(synth.splitmix64(1012, 1)[0])
    assert m[0, 0] == 2.0 * ((u >> 11) * 2.0 ** -53) - 1.0
    assert np.all(m >= -1.0) and np.all(m < 1.0)


def test_scene_is_the_rigid_image_plus_noise():
    n = 1 << 14
    m, p, perm = synth.pair(n, 1014, 2014, 3014)
    assert sorted(perm.tolist()) == list(range(n)) and not np.array_equal(perm, np.arange(n))
    g = (p - (m[perm] @ synth.rotation(5.0).T + np.array([0.05, -0.03, 0.02]))) / 1e-3
    assert abs(g.mean()) < 0.03 and abs(g.std() - 1.0) < 0.03 and np.abs(g).max() < 6.0
    assert abs(np.corrcoef(g[:, 0], g[:, 1])[0, 1]) < 0.03
    R = synth.rotation(5.0)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-15) and np.allclose(R @ np.array([1.0, 2.0, 3.0]), [1.0, 2.0, 3.0], atol=1e-14)


def test_config5_pairs_differ_and_vary_the_angle():
    a, _ = synth.config5_pair(0, 256)
    b, _ = synth.config5_pair(1, 256)
    assert not np.array_equal(a, b)
    angles = {2.0 + 6.0 * ((k * 37) % 101) / 100.0 for k in range(1024)}
    assert min(angles) == 2.0 and max(angles) == 8.0
