"""GPU parity tests (run on the B200 box): the CUDA path, called through the C-ABI, against the CPU
oracle on the same inputs and against the committed golden vectors.

Bars (BASELINE.json north_star): correspondence indices bit-exact; s, R, t and the error value within
1e-5 relative; identical iteration count.
"""
import json
import os
import zlib

import numpy as np
import pytest

from conftest import GOLDEN, synthetic_pair
from oracle import oracle as O

pytestmark = pytest.mark.gpu

RTOL = 1e-5   # north_star tolerance for floating-point results


@pytest.fixture(scope="module")
def icp():
    import icp_b200
    icp_b200.load_library()
    return icp_b200


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


# ---- K1: closest_matrix -------------------------------------------------------------------------

@pytest.mark.parametrize("mname,pname", [("cow_ref", "cow_tr1"), ("cow_ref", "cow_tr2"), ("horse_ref", "horse_tr1"),
                                         ("horse_ref", "horse_tr2"), ("bun000", "bun045"), ("bun045", "bun000")])
def test_closest_matrix_indices_bit_exact(icp, clouds, mname, pname):
    m, p = clouds(mname), clouds(pname)
    idx_o, Y_o = O.closest_matrix(p, m, O.NN_MULSQRT)
    Y, idx = icp.closest_matrix(p, m, return_index=True)
    assert np.array_equal(idx, idx_o)
    assert np.array_equal(Y, Y_o)          # Y columns are copies of model columns: exact


def test_closest_small_and_ties(icp):
    m = np.array([[1.0, 0, 0], [-1.0, 0, 0], [1.0, 0, 0], [0, 5.0, 0]])
    p = np.array([[0.0, 0, 0], [0.9, 0, 0], [0, 4.0, 0]])
    Y, idx = icp.closest_matrix(p, m, return_index=True)
    assert idx.tolist() == [0, 0, 3]       # first minimum wins (Eigen minCoeff, src/cpu.cc:21-22)
    assert icp.compute_distance_w_naive(m, [0.0, 0.1, 0]) == 0
    # single model point / single scene point
    assert icp.closest_matrix(p, m[:1], return_index=True)[1].tolist() == [0, 0, 0]


def test_closest_duplicates_across_chunks(icp):
    # exact duplicates thousands of indices apart: the FP32 filter cannot separate them, the FP64
    # re-rank must return the lowest index
    rng = np.random.default_rng(5)
    base = rng.uniform(-1, 1, (3000, 3))
    m = np.concatenate([base, base[::-1], base])          # every point three times
    p = base[rng.permutation(3000)[:777]] + 1e-9
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    _, idx = icp.closest_matrix(p, m, return_index=True)
    assert np.array_equal(idx, idx_o)
    assert (idx < 3000).all()


def test_closest_near_ties_below_fp32_resolution(icp):
    # candidates whose distances differ by ~1e-12 relative, on a cloud with a large common offset
    rng = np.random.default_rng(7)
    p = rng.uniform(-1, 1, (2048, 3)) + 100.0
    d = rng.standard_normal((2048, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = 0.01
    m = np.concatenate([p + r * d * (1 + 1e-12), p - r * d, p + r * d[:, [1, 2, 0]] * (1 + 3e-12)])
    m = m[rng.permutation(len(m))]
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    _, idx = icp.closest_matrix(p, m, return_index=True)
    assert np.array_equal(idx, idx_o)


@pytest.mark.parametrize("form", ["auto", "inline", "kernel"])
def test_closest_exact_scan_paths(icp, monkeypatch, form):
    # nn_exact work decomposition: (a) a handful of undecidable points against a model long enough to be cut
    # into several segments (ragged last batch, cross-segment merge by ticket); (b) every scene point
    # undecidable (one segment, many batches); (c) a scene far away from the model (large tau).
    # ICP_EXACT (read at icp_create, and at every use by the hidden context of the one-shot wrappers) forces the rescan
    # inside nn_refine or the separate nn_exact kernel, so both forms meet every case whatever the automatic choice is
    monkeypatch.setenv("ICP_EXACT", form)
    rng = np.random.default_rng(21)
    base = rng.uniform(-1, 1, (20000, 3))
    m = np.concatenate([base, base[:50][::-1]])           # 50 points exist twice, 20 000 indices apart
    p = base[[3, 11, 17, 23, 29, 31, 47]] + 1e-10
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    _, idx = icp.closest_matrix(p, m, return_index=True)
    assert np.array_equal(idx, idx_o) and (idx < 50).all()
    m2 = np.concatenate([base[:4500], base[:4500]])       # everything twice
    p2 = base[rng.permutation(4500)[:1501]] - 1e-10
    idx_o, _ = O.closest_matrix(p2, m2, O.NN_MULSQRT)
    ctx = icp.Context(0)
    ctx.set_model(m2)
    ctx.set_scene(p2)
    idx, _ = ctx.closest(want_Y=False)
    assert np.array_equal(idx, idx_o) and (idx < 4500).all()
    assert ctx.stats()["rerank_full"] == len(p2)
    ctx.close()
    p3 = rng.uniform(-1, 1, (3000, 3)) + np.array([40.0, -25.0, 60.0])
    idx_o, _ = O.closest_matrix(p3, base, O.NN_MULSQRT)
    _, idx = icp.closest_matrix(p3, base, return_index=True)
    assert np.array_equal(idx, idx_o)


@pytest.mark.parametrize("form", ["auto", "inline", "kernel"])
def test_closest_exact_scan_large_model(icp, monkeypatch, form):
    # a model cut into many slots per scene tile: (a) a few duplicates far apart -> two distant slot ranges qualify for
    # each undecided point; (b) every scene point undecided
    monkeypatch.setenv("ICP_EXACT", form)
    rng = np.random.default_rng(23)
    base = rng.uniform(-1, 1, (140000, 3))
    m = np.concatenate([base, base[:50][::-1]])
    p = base[[3, 11, 17, 23, 29, 31, 47]] + 1e-10
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    _, idx = icp.closest_matrix(p, m, return_index=True)
    assert np.array_equal(idx, idx_o) and (idx < 50).all()
    m2 = np.concatenate([base[:70000], base[:70000]])
    p2 = base[rng.permutation(70000)[:1501]] - 1e-10
    idx_o, _ = O.closest_matrix(p2, m2, O.NN_MULSQRT)
    ctx = icp.Context(0)
    ctx.set_model(m2)
    ctx.set_scene(p2)
    idx, _ = ctx.closest(want_Y=False)
    assert np.array_equal(idx, idx_o) and (idx < 70000).all()
    assert ctx.stats()["rerank_full"] == len(p2)
    ctx.close()


@pytest.mark.parametrize("nn", ["brute", "tree"])
def test_non_finite_input_follows_the_reference(icp, nn):
    # Eigen's minCoeff visitor starts at column 0 and only moves on a strict `<` (src/cpu.cc:21-22): a scene point with
    # a NaN or infinite coordinate has only NaN / +inf distances and matches index 0; a NaN model point is never
    # chosen, except model point 0, which is then never left.  The oracle restates exactly that.  None of it may fault.
    rng = np.random.default_rng(31)
    m = rng.uniform(-1, 1, (5000, 3))
    p = rng.uniform(-1, 1, (3000, 3))
    cases = []
    p1 = p.copy(); p1[7, 1] = np.nan; p1[1999, 0] = np.inf; p1[2999] = -np.inf; p1[100, 2] = 1e300
    cases.append((m, p1))
    m2 = m.copy(); m2[2500, 0] = np.nan; m2[4999] = np.nan                      # mid-array and last
    cases.append((m2, p))
    m3 = m.copy(); m3[0, 2] = np.nan                                            # index 0: every match is 0
    cases.append((m3, p))
    for mm, pp in cases:
        idx_o, _ = O.closest_matrix(pp, mm, O.NN_MULSQRT)
        ctx = icp.Context(0)
        ctx.set_nn_mode(nn)
        ctx.set_model(mm)
        ctx.set_scene(pp)
        idx, _ = ctx.closest(want_Y=False)
        assert np.array_equal(idx, idx_o)
        it, hist = ctx.run(3, 1e-5, False)           # the loop must survive it too (NaN transform, like the reference)
        assert it == 3 and (ctx.indices() >= 0).all() and (ctx.indices() < len(mm)).all()
        ctx.close()
    assert (O.closest_matrix(p1, m, O.NN_MULSQRT)[0][[7, 1999, 2999, 100]] == 0).all()
    assert (O.closest_matrix(p, m3, O.NN_MULSQRT)[0] == 0).all()


def test_fused_alignment_equals_split_kernels(icp, clouds, monkeypatch):
    # the cooperative K2+K3+K4 kernel and the separate moments / solve / transform kernels are the same maths
    m, p = clouds("cow_ref"), clouds("cow_tr1")
    res = []
    for mode in ("fused", "split"):
        monkeypatch.setenv("ICP_ALIGN", mode)
        ctx = icp.Context(0)
        ctx.set_model(m)
        ctx.set_scene(p)
        it, hist = ctx.run(30)
        res.append((it, hist, ctx.get_scene(), ctx.transform(), ctx.stats()["kernels_launched"]))
        ctx.close()
    assert res[0][0] == res[1][0] == 7
    assert res[0][4] < res[1][4]                           # 3 launches per iteration instead of 4
    assert np.allclose(res[0][1], res[1][1], rtol=1e-9, atol=1e-18)
    assert np.abs(res[0][2] - res[1][2]).max() < 1e-12


def test_batch_run_matches_oracle_per_pair(icp):
    # BASELINE config 5 in miniature: independent pairs through icp_batch_run, 3 concurrent lanes
    models, scenes = [], []
    for k in range(7):
        m, p, _ = synthetic_pair(1500 + 211 * k, seed=100 + k, angle_deg=2.0 + k, noise=1e-3 if k % 2 else 0.0)
        models.append(m)
        scenes.append(p)
    res = icp.batch_run(models, scenes, max_iter=25, lanes=3)
    for k in range(7):
        ro = O.icp_run(models[k], scenes[k], 25)
        assert res["iters"][k] == ro["iters"], k
        assert abs(res["err"][k] - ro["err"][-1]) <= 1e-5 * abs(ro["err"][-1]) + 1e-18
        assert np.abs(res["new_p"][k] - ro["new_p"]).max() < 1e-9
    # error path: a size mismatch in one pair fails the batch like the reference's exit(-1)
    with pytest.raises(icp.IcpError):
        icp.batch_run([models[0]], [scenes[1]], max_iter=5)


def test_batch_run_full_size_pair_matches_oracle(icp):
    # BASELINE config 5 at its stated point count: one 65 536-point noisy pair of the batch (SURVEY 8d recipe, pair 3)
    # through icp_batch_run, next to a small pair so that two lanes are busy; iteration count, error history end,
    # transform and new_p against the oracle's full loop (a few seconds on the host's cores)
    from icp_b200 import synth
    m, p = synth.config5_pair(3)
    ms, ps, _ = synthetic_pair(3000, seed=77)
    res = icp.batch_run([m, ms], [p, ps], max_iter=30, lanes=2)
    ro = O.icp_run(m, p, 30)
    assert res["iters"][0] == ro["iters"]
    assert abs(res["err"][0] - ro["err"][-1]) <= RTOL * abs(ro["err"][-1])
    assert np.abs(res["new_p"][0] - ro["new_p"]).max() < 1e-9
    s_, R_, t_ = res["srt"][0][0], res["srt"][0][1:10].reshape(3, 3), res["srt"][0][10:13]
    assert abs(s_ - ro["s"][-1]) <= RTOL * abs(ro["s"][-1]) and rel(R_, ro["R"][-1]) <= RTOL and rel(t_, ro["t"][-1]) <= RTOL
    assert res["kernels_launched"] >= 3 * ro["iters"]


@pytest.mark.parametrize("n,seed", [(1000, 0), (5000, 1), (65536, 2)])
def test_closest_synthetic(icp, n, seed):
    m, p, _ = synthetic_pair(n, seed=seed)
    idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
    _, idx = icp.closest_matrix(p, m, return_index=True)
    assert np.array_equal(idx, idx_o)


def test_closest_ragged_sizes(icp):
    rng = np.random.default_rng(11)
    for np_, nm in [(1, 1), (3, 7), (1025, 511), (1023, 513), (2049, 64), (17, 4097)]:
        p, m = rng.standard_normal((np_, 3)), rng.standard_normal((nm, 3))
        idx_o, _ = O.closest_matrix(p, m, O.NN_MULSQRT)
        _, idx = icp.closest_matrix(p, m, return_index=True)
        assert np.array_equal(idx, idx_o), (np_, nm)


def test_closest_large_scale_property(icp):
    # 2^20 points: too many for the oracle; the scene is the model itself, permuted -> idx must be the
    # inverse permutation exactly (distance 0 is the unique minimum)
    rng = np.random.default_rng(3)
    n = 1 << 20
    m = rng.uniform(-1, 1, (n, 3))
    perm = rng.permutation(n)
    ctx = icp.Context(0)
    ctx.set_model(m)
    ctx.set_scene(m[perm])
    idx, _ = ctx.closest(want_Y=False)
    assert np.array_equal(idx, perm.astype(np.int32))
    st = ctx.stats()
    assert st["nn_pairs"] == n * n
    ctx.close()


def test_closest_full_size_property_4m(icp):
    # BASELINE's largest single cloud (2^22 points, 1.76e13 pairs): the scene is a rigidly moved copy of the
    # model in a different order, with a shift far below the point spacing, so the exact answer is the inverse
    # permutation; then one alignment step must recover the motion (s = 1, R, t) to 1e-9.
    rng = np.random.default_rng(22)
    n = 1 << 22
    m = rng.uniform(-1, 1, (n, 3))
    perm = rng.permutation(n)
    a = 1e-7
    R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1.0]])
    t = np.array([2e-7, -1e-7, 3e-7])
    ctx = icp.Context(0)
    ctx.set_model(m)
    ctx.set_scene((m @ R.T + t)[perm])
    idx, _ = ctx.closest(want_Y=False)
    assert np.array_equal(idx, perm.astype(np.int32))
    fa = ctx.find_alignment()
    assert abs(fa["s"] - 1.0) < 1e-9 and np.abs(fa["R"] - R.T).max() < 1e-9 and np.abs(fa["t"] + R.T @ t).max() < 1e-9
    assert fa["err"] < 1e-18 * n
    ctx.close()


# ---- K2-K4 operators ------------------------------------------------------------------------------

def test_find_alignment_and_err_compute(icp, clouds):
    m, p = clouds("cow_ref"), clouds("cow_tr1")
    ctx = icp.Context(0)
    ctx.set_model(m)
    ctx.set_scene(p)
    idx, Y = ctx.closest()
    ro = O.find_alignment(p, Y)
    mu_p, mu_y = ctx.centroid()
    op, oy = O.centroid(p, Y)
    assert rel(mu_p, op) < 1e-12 and rel(mu_y, oy) < 1e-12
    r = ctx.find_alignment()
    assert abs(r["s"] - ro["s"]) / ro["s"] < RTOL
    assert rel(r["R"], ro["R"]) < RTOL and rel(r["t"], ro["t"]) < RTOL
    assert abs(r["err"] - ro["err"]) / ro["err"] < RTOL
    # err_compute_alignment does not move the scene; err_compute does
    assert np.array_equal(ctx.get_scene(), p)
    e_o, newp_o = O.err_compute(p, ro["s"], ro["R"], ro["t"], Y)
    e = ctx.err_compute(in_place=True)
    assert abs(e - e_o) / e_o < RTOL
    assert rel(ctx.get_scene(), newp_o) < 1e-12
    ctx.close()


def test_reference_shaped_wrappers(icp, clouds):
    m, p = clouds("cow_ref"), clouds("cow_tr2")
    idx, Y = O.closest_matrix(p, m)
    ro = O.find_alignment(p, Y)
    sr = ro["s"] * ro["R"]
    e_o, newp_o = O.err_compute(p, ro["s"], ro["R"], ro["t"], Y)
    e, newp = icp.compute_err_w(Y, p, True, sr, ro["t"])
    assert abs(e - e_o) / e_o < RTOL and rel(newp, newp_o) < 1e-12
    e2, same = icp.compute_err_w(Y, p, False, sr, ro["t"])
    assert abs(e2 - O.err_compute_alignment(p, ro["s"], ro["R"], ro["t"], Y)) / e_o < RTOL
    assert np.array_equal(same, p)
    assert abs(icp.err_compute_alignment(p, ro["s"], ro["R"], ro["t"], Y) - e2) < 1e-9 * e2
    mu_p, mu_y, pp, yp = O.centroid(p, Y, centred=True)
    assert rel(icp.substract_col_w(p, mu_p), pp) < 1e-15
    d_caps, sp = icp.y_p_norm_w(yp, pp)
    assert abs(d_caps - (yp * yp).sum()) / (yp * yp).sum() < 1e-10
    assert abs(sp - (pp * pp).sum()) / (pp * pp).sum() < 1e-10


# ---- the loop ---------------------------------------------------------------------------------------

@pytest.mark.parametrize("case", ["cow_tr1", "cow_tr2", "horse_tr1", "horse_tr2", "bunny"])
def test_full_loop_matches_golden(icp, clouds, golden, case):
    meta = json.load(open(os.path.join(GOLDEN, "meta.json")))["cases"][case]
    g = golden(case)
    m, p = clouds(meta["model"]), clouds(meta["scene"])
    ctx = icp.Context(0)
    ctx.set_model(m)
    ctx.set_scene(p)
    iters, err = ctx.run(meta["max_iter"], 1e-5, check_sizes=not meta["allow_unequal"])
    assert iters == int(g["iters"])                       # identical iteration count
    assert rel(err, g["err"]) < RTOL and np.all(np.abs(err - g["err"]) <= RTOL * np.abs(g["err"]))
    assert np.array_equal(ctx.indices(), g["idx_last"])   # correspondences of the last iteration
    s, R, t = ctx.transform()
    assert abs(s - g["s"][-1]) / g["s"][-1] < RTOL and rel(R, g["R"][-1]) < RTOL
    assert np.abs(t - g["t"][-1]).max() < RTOL * max(np.abs(g["t"][-1]).max(), np.abs(p).max() * 1e-3)
    assert rel(ctx.get_scene(), g["new_p"]) < RTOL
    ctx.close()


@pytest.mark.parametrize("case", ["cow_tr1", "horse_tr2"])
def test_per_iteration_indices_match_golden_crc(icp, clouds, golden, case):
    # step the loop operator by operator (closest -> find_alignment -> err_compute) and compare the
    # correspondence indices of EVERY iteration with the oracle's
    meta = json.load(open(os.path.join(GOLDEN, "meta.json")))["cases"][case]
    g = golden(case)
    ctx = icp.Context(0)
    ctx.set_model(clouds(meta["model"]))
    ctx.set_scene(clouds(meta["scene"]))
    n = len(clouds(meta["scene"]))
    for i in range(int(g["iters"])):
        idx, _ = ctx.closest(want_Y=False)
        assert zlib.crc32(idx.tobytes()) == int(g["idx_crc"][i]), f"iteration {i}"
        r = ctx.find_alignment()
        err = (r["err"] + ctx.err_compute(in_place=True)) / n
        assert abs(err - g["err"][i]) <= RTOL * g["err"][i]
    ctx.close()


def test_icp_class_mirrors_reference_behaviour(icp, clouds, capsys):
    m, p = clouds("cow_ref"), clouds("cow_tr1")
    obj = icp.ICP(m, p, 10)
    assert obj.find_corresponding() == 7
    err = capsys.readouterr().err.splitlines()
    assert err[0].startswith("[ICP] iteration number 0 | error value = 0.0224295")
    assert len(err) == 7
    assert obj.new_p.shape == p.shape
    with pytest.raises(SystemExit):
        icp.ICP(m, p[:100], 3).find_corresponding()
    assert "same number of points" in capsys.readouterr().err
    with pytest.raises(SystemExit):
        icp.ICP(m[:3], p[:3], 3).find_corresponding()
    # reset + rerun gives identical results (device-side backup of the scene)
    obj.ctx.reset_scene()
    it2, hist2 = obj.ctx.run(10)
    assert it2 == 7 and np.array_equal(hist2, obj.errors)


# ---- K3 robustness: fast quartic/adjugate path and its Jacobi fallback -------------------------------

@pytest.mark.parametrize("kind", ["generic", "identity", "planar", "collinear", "scaled", "big_rotation", "tiny", "far_from_origin"])
def test_horn_solver_cases(icp, kind):
    rng = np.random.default_rng(42)
    n = 500
    y = rng.standard_normal((n, 3))
    a = {"big_rotation": 2.9}.get(kind, 0.02)
    ax = np.array([0.3, -0.5, 0.81]); ax /= np.linalg.norm(ax)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    R = np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * (K @ K)
    s, t = 1.0, np.array([0.01, -0.02, 0.005])
    if kind == "identity":
        R, t = np.eye(3), np.zeros(3)
    if kind == "planar":
        y[:, 2] = 0.0
    if kind == "collinear":                   # rotation about the line is undetermined: only the residual is checked
        y = np.outer(rng.standard_normal(n), [1.0, 2.0, -0.5])
    if kind == "scaled":
        s = 1.02
    if kind == "tiny":
        y *= 1e-4; t *= 1e-4
    if kind == "far_from_origin":
        y += 1000.0
    # scene = inverse transform of the model, so that the alignment p -> y is (s, R, t)
    p = ((y - t) / s) @ R                     # p_j = R^T (y_j - t) / s
    r = icp.find_alignment(p, y)          # K2 + K3 + K4 on explicit pairs
    ro = O.find_alignment(p, y)
    assert abs(r["s"] - ro["s"]) < 1e-9 * ro["s"]
    assert r["err"] <= 1e-18 * n * max(1.0, np.abs(y).max() ** 2) + 1e-9 * abs(ro["err"])
    if kind == "collinear":
        return
    assert np.abs(r["R"] - ro["R"]).max() < 1e-8
    assert np.abs(r["t"] - ro["t"]).max() < 1e-8 * max(1.0, np.abs(y).max())
    assert np.abs(r["R"] - R).max() < 1e-7 and abs(r["s"] - s) < 1e-9
