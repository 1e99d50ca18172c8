"""Compact workload for compute-sanitizer (one tool per gpurun call): every kernel of the hot path on small inputs,
including ragged sizes, the exact-scan paths, a full loop, the one-shot wrappers and a two-lane batch."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "iterative-closest-point_b200")); sys.path.insert(0, ROOT)
import icp_b200

rng = np.random.default_rng(1)
for np_, nm in [(1, 1), (3, 7), (1025, 511), (1023, 513), (2049, 64), (17, 4097), (3000, 9000)]:
    p, m = rng.standard_normal((np_, 3)), rng.standard_normal((nm, 3))
    Y, idx = icp_b200.closest_matrix(p, m, return_index=True)
    d = np.linalg.norm(p[:, None, :] - m[None, :, :], axis=2) if np_ * nm < 3e7 else None
    if d is not None:
        assert np.array_equal(idx, d.argmin(1)), (np_, nm)
base = rng.uniform(-1, 1, (6000, 3))
m = np.concatenate([base, base[:40][::-1]])
p = base[[3, 11, 17, 23, 29, 31, 39]] + 1e-10
_, idx = icp_b200.closest_matrix(p, m, return_index=True)
assert (idx < 40).all()
m2 = np.concatenate([base[:1500], base[:1500]])
_, idx = icp_b200.closest_matrix(base[:700] - 1e-10, m2, return_index=True)
assert (idx < 1500).all()
a = np.deg2rad(4.0)
R = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
mm = rng.uniform(-1, 1, (2500, 3))
pp = (mm @ R.T + 0.02)[rng.permutation(2500)]
for mode in ("split", "fused"):
    os.environ["ICP_ALIGN"] = mode
    ctx = icp_b200.Context(0)
    ctx.set_model(mm); ctx.set_scene(pp)
    it, hist = ctx.run(15)
    ctx.find_alignment(); ctx.err_compute(False); ctx.centroid(); ctx.get_scene(); ctx.indices()
    ctx.close()
    print(mode, "iterations", it, "err", hist[-1])
res = icp_b200.batch_run([mm, mm[:2000]], [pp, pp[:2000]], max_iter=10, lanes=2, check_sizes=False)
print("batch", res["iters"])
icp_b200.find_alignment(pp, mm); icp_b200.substract_col_w(pp, pp.mean(0)); icp_b200.y_p_norm_w(mm, pp)
print("sanitize_case ok")
