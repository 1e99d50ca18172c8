"""GPU parity tests of the exact tree search (ICP_NN_TREE): same bar as the brute-force path —
correspondence indices bit-exact against the oracle / the golden vectors — plus tree == brute force on inputs the
oracle cannot finish."""
import json
import os
import zlib

import numpy as np
import pytest

from conftest import GOLDEN, synthetic_pair
from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def icp():
    import icp_b200
    icp_b200.load_library()
    return icp_b200


def tree_closest(icp, p, m, expect_mode="tree"):
    ctx = icp.Context(0)
    ctx.set_nn_mode("tree")
    ctx.set_model(m)
    ctx.set_scene(p)
    assert ctx.nn_mode() == expect_mode
    idx, Y = ctx.closest()
    st = ctx.stats()
    ctx.close()
    return idx, Y, st


@pytest.mark.parametrize("mname,pname", [("cow_ref", "cow_tr1"), ("cow_ref", "cow_tr2"), ("horse_ref", "horse_tr1"),
                                         ("horse_ref", "horse_tr2"), ("bun000", "bun045"), ("bun045", "bun000")])
def test_tree_closest_bit_exact(icp, clouds, mname, pname):
    m, p = clouds(mname), clouds(pname)
    idx_o, Y_o = O.closest_matrix(p, m, O.NN_MULSQRT)
    idx, Y, st = tree_closest(icp, p, m)
    assert np.array_equal(idx, idx_o)
    assert np.array_equal(Y, Y_o)
    assert st["kernels_launched"] == 1          # one search kernel instead of filter + refine + exact


def test_tree_ties_duplicates_and_near_ties(icp):
    m = np.array([[1.0, 0, 0], [-1.0, 0, 0], [1.0, 0, 0], [0, 5.0, 0]])
    p = np.array([[0.0, 0, 0], [0.9, 0, 0], [0, 4.0, 0]])
    assert tree_closest(icp, p, m)[0].tolist() == [0, 0, 3]          # first minimum (src/cpu.cc:21-22)
    assert tree_closest(icp, p, m[:1])[0].tolist() == [0, 0, 0]
    rng = np.random.default_rng(5)
    base = rng.uniform(-1, 1, (3000, 3))
    m = np.concatenate([base, base[::-1], base])                     # every point three times
    p = base[rng.permutation(3000)[:777]] + 1e-9
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    idx = tree_closest(icp, p, m)[0]
    assert np.array_equal(idx, idx_o) and (idx < 3000).all()
    # distances that differ by ~1e-12 relative, on a cloud with a large common offset
    rng = np.random.default_rng(7)
    p = rng.uniform(-1, 1, (2048, 3)) + 100.0
    d = rng.standard_normal((2048, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    m = np.concatenate([p + 0.01 * d * (1 + 1e-12), p - 0.01 * d, p + 0.01 * d[:, [1, 2, 0]] * (1 + 3e-12)])
    m = m[rng.permutation(len(m))]
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    assert np.array_equal(tree_closest(icp, p, m)[0], idx_o)


def test_tree_ragged_far_and_degenerate(icp):
    rng = np.random.default_rng(11)
    for np_, nm in [(1, 1), (3, 7), (1025, 511), (1023, 513), (2049, 64), (17, 4097), (5, 8), (5, 9)]:
        p, m = rng.standard_normal((np_, 3)), rng.standard_normal((nm, 3))
        idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
        assert np.array_equal(tree_closest(icp, p, m)[0], idx_o), (np_, nm)
    base = rng.uniform(-1, 1, (20000, 3))
    p3 = rng.uniform(-1, 1, (3000, 3)) + np.array([40.0, -25.0, 60.0])       # scene far outside the model's box
    idx_o, _ = O.closest_matrix(p3, base, O.NN_MULSQRT)
    assert np.array_equal(tree_closest(icp, p3, base)[0], idx_o)
    flat = base.copy()
    flat[:, 2] = 0.25                                                         # planar model, collinear scene
    line = np.stack([np.linspace(-2, 2, 999), np.zeros(999), np.zeros(999)], 1)
    idx_o, _ = O.closest_matrix(line, flat, O.NN_MULSQRT)
    assert np.array_equal(tree_closest(icp, line, flat)[0], idx_o)
    huge = np.array([[1e25, 0, 0], [0, -3e22, 1e21], [1e20, 1e20, 1e20]])    # FP32 squares overflow: FP64 scan fallback
    idx_o, _ = O.closest_matrix(huge, base[:700], O.NN_MULSQRT)
    assert np.array_equal(tree_closest(icp, huge, base[:700])[0], idx_o)
    same = np.tile(np.array([[0.3, -0.2, 0.1]]), (100, 1))                    # all model points identical
    assert tree_closest(icp, p3[:10], same)[0].tolist() == [0] * 10


def test_tree_scale_outside_range_keeps_brute_force(icp):
    rng = np.random.default_rng(13)
    m, p = rng.uniform(-1, 1, (2000, 3)) * 1e-15, rng.uniform(-1, 1, (500, 3)) * 1e-15
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    idx, _, st = tree_closest(icp, p, m, expect_mode="brute")
    assert np.array_equal(idx, idx_o) and st["kernels_launched"] == 2      # filter + refine (undecided points rescanned in place)


@pytest.mark.parametrize("n,seed", [(5000, 1), (65536, 2)])
def test_tree_synthetic(icp, n, seed):
    m, p, _ = synthetic_pair(n, seed=seed)
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    assert np.array_equal(tree_closest(icp, p, m)[0], idx_o)


@pytest.mark.parametrize("case", ["cow_tr1", "cow_tr2", "horse_tr1", "horse_tr2", "bunny"])
def test_tree_full_loop_matches_golden_and_brute(icp, clouds, golden, case):
    meta = json.load(open(os.path.join(GOLDEN, "meta.json")))["cases"][case]
    g = golden(case)
    m, p = clouds(meta["model"]), clouds(meta["scene"])
    out = {}
    for mode in ("brute", "tree"):
        ctx = icp.Context(0)
        ctx.set_nn_mode(mode)
        ctx.set_model(m)
        ctx.set_scene(p)
        iters, err = ctx.run(meta["max_iter"], 1e-5, check_sizes=not meta["allow_unequal"])
        out[mode] = (iters, err, ctx.indices(), ctx.get_scene(), ctx.stats()["kernels_launched"])
        ctx.close()
    it_t, err_t, idx_t, sc_t, launches_t = out["tree"]
    assert it_t == int(g["iters"])
    assert np.array_equal(idx_t, g["idx_last"])
    # identical correspondences feed identical alignment kernels: everything downstream is bit-identical
    assert it_t == out["brute"][0] and np.array_equal(err_t, out["brute"][1]) and np.array_equal(sc_t, out["brute"][3])
    assert launches_t < out["brute"][4]


@pytest.mark.parametrize("case", ["cow_tr1", "horse_tr2"])
def test_tree_per_iteration_indices_match_golden_crc(icp, clouds, golden, case):
    # operator by operator; from the second iteration on the search starts from the previous matches
    meta = json.load(open(os.path.join(GOLDEN, "meta.json")))["cases"][case]
    g = golden(case)
    ctx = icp.Context(0)
    ctx.set_nn_mode("tree")
    ctx.set_model(clouds(meta["model"]))
    ctx.set_scene(clouds(meta["scene"]))
    for i in range(int(g["iters"])):
        idx, _ = ctx.closest(want_Y=False)
        assert zlib.crc32(idx.tobytes()) == int(g["idx_crc"][i]), f"iteration {i}"
        ctx.find_alignment()
        ctx.err_compute(in_place=True)
    ctx.close()


def test_tree_equals_brute_force_on_noisy_large_clouds(icp):
    # 2^18 noisy synthetic pair (the oracle would need minutes): both searches on the same context
    m, p, _ = synthetic_pair(1 << 18, seed=4)
    ctx = icp.Context(0)
    ctx.set_model(m)
    ctx.set_scene(p)
    idx_b, _ = ctx.closest(want_Y=False)
    ctx.set_nn_mode("tree")
    idx_t, _ = ctx.closest(want_Y=False)         # starts from the brute-force matches (use_prev)
    assert ctx.nn_mode() == "tree" and np.array_equal(idx_t, idx_b)
    ctx.reset_scene()                            # forget the matches: cold search
    idx_c, _ = ctx.closest(want_Y=False)
    assert np.array_equal(idx_c, idx_b)
    # stale matches as the starting point: shuffle the scene's matches by moving the scene
    ctx.find_alignment()
    ctx.err_compute(in_place=True)
    idx_t2, _ = ctx.closest(want_Y=False)
    ctx.set_nn_mode("brute")
    idx_b2, _ = ctx.closest(want_Y=False)
    assert np.array_equal(idx_t2, idx_b2)
    ctx.close()


def test_tree_large_model_small_scene(icp):
    # a model large enough for the global k-d levels of the build, searched by the warp-per-point form (small scene)
    # and, with the same model, by the thread-per-point form (forced): both equal brute force
    m, p, _ = synthetic_pair(1 << 18, seed=9)
    p = np.ascontiguousarray(p[::37])
    ctx = icp.Context(0)
    ctx.set_model(m)
    ctx.set_scene(p)
    idx_b, _ = ctx.closest(want_Y=False)
    ctx.set_nn_mode("tree")
    ctx.reset_scene()
    idx_t, _ = ctx.closest(want_Y=False)
    assert ctx.nn_mode() == "tree" and np.array_equal(idx_t, idx_b)
    ctx.close()


def test_tree_in_hidden_contexts_follows_environment(icp, clouds, monkeypatch):
    # one-shot wrappers and icp_batch_run lanes keep their contexts between calls: ICP_NN is read at every use
    m, p = clouds("cow_ref"), clouds("cow_tr1")
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    models, scenes = [], []
    for k in range(5):
        mk, pk, _ = synthetic_pair(1500 + 211 * k, seed=100 + k, angle_deg=2.0 + k, noise=1e-3 if k % 2 else 0.0)
        models.append(mk)
        scenes.append(pk)
    res = {}
    for nn in ("brute", "tree", "brute"):
        monkeypatch.setenv("ICP_NN", nn)
        assert np.array_equal(icp.closest_matrix(p, m, return_index=True)[1], idx_o)
        res[nn] = icp.batch_run(models, scenes, max_iter=25, lanes=3)
    assert np.array_equal(res["tree"]["iters"], res["brute"]["iters"])
    assert np.array_equal(res["tree"]["err"], res["brute"]["err"])
    for a, b in zip(res["tree"]["new_p"], res["brute"]["new_p"]):
        assert np.array_equal(a, b)


def test_tree_full_size_property_4m(icp):
    # BASELINE's largest cloud: scene = rigidly moved, permuted copy of the model with a shift far below the point
    # spacing, so the exact answer is the inverse permutation
    rng = np.random.default_rng(22)
    n = 1 << 22
    m = rng.uniform(-1, 1, (n, 3))
    perm = rng.permutation(n)
    a = 1e-7
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1.0]])
    ctx = icp.Context(0)
    ctx.set_nn_mode("tree")
    ctx.set_model(m)
    ctx.set_scene((m @ R.T + np.array([2e-7, -1e-7, 3e-7]))[perm])
    idx, _ = ctx.closest(want_Y=False)
    assert np.array_equal(idx, perm.astype(np.int32))
    ctx.close()


def test_auto_mode_takes_the_tree_only_inside_loops_on_small_clouds(icp, clouds, golden, monkeypatch):
    # default mode (ICP_NN unset): icp_run on clouds too small to fill one wave of the brute-force filter searches through
    # the tree - same indices, iteration count and error history as the golden vectors; a one-shot icp_closest and any
    # cloud above the limit stay brute force
    monkeypatch.delenv("ICP_NN", raising=False)
    g = golden("cow_tr1")
    ctx = icp.Context(0)
    ctx.set_model(clouds("cow_ref"))
    ctx.set_scene(clouds("cow_tr1"))
    idx, _ = ctx.closest(want_Y=False)
    assert ctx.nn_mode() == "brute" and np.array_equal(idx, g["idx_first"])
    ctx.reset_scene()
    it, hist = ctx.run(30)
    assert ctx.nn_mode() == "tree"
    assert it == int(g["iters"]) and np.all(np.abs(hist - g["err"]) <= 1e-5 * np.abs(g["err"]))
    assert np.array_equal(ctx.indices(), g["idx_last"]) and np.abs(ctx.get_scene() - g["new_p"]).max() < 1e-9
    ctx.reset_scene()
    it1, _ = ctx.run(1)                                   # a single iteration does not pay for the build
    assert it1 == 1 and ctx.nn_mode() == "brute"
    ctx.set_nn_mode("brute")
    ctx.reset_scene()
    it_b, hist_b = ctx.run(30)
    assert ctx.nn_mode() == "brute" and it_b == it and np.array_equal(hist_b, hist)      # everything after idx is shared code
    ctx.set_nn_mode("auto")
    g2 = golden("horse_tr2")
    ctx.set_model(clouds("horse_ref"))
    ctx.set_scene(clouds("horse_tr2"))
    it2, _ = ctx.run(30)
    assert ctx.nn_mode() == "brute" and it2 == int(g2["iters"]) and np.array_equal(ctx.indices(), g2["idx_last"])
    ctx.close()
