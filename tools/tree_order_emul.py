"""CPU emulation behind profiles/r01i_tree_order.md: how many boxes / leaves of the tree search's hierarchy lie inside
a search's final radius, as a function of the ORDER of the model points (Morton runs, median splits inside blocks of B
consecutive Morton-sorted points, k-d order from the top).  numpy + scipy only; no GPU, no oracle.

  python tools/tree_order_emul.py wide      # horse_ref vs the converged horse_tr1 scene, 32-point leaves, 32-ary blocks
  python tools/tree_order_emul.py binary    # synthetic 2^17 uniform cloud, 8-point leaves, binary heap
"""
import lzma
import os
import sys

import numpy as np
from scipy.spatial import cKDTree

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import GOLDEN, synthetic_pair  # noqa: E402


def cloud(name):
    rows = lzma.open(os.path.join(GOLDEN, "clouds", name + ".txt.xz"), "rt").read().splitlines()[1:]
    return np.array([[float(x) for x in r.split(",")[:3]] for r in rows])


def morton_order(x, bits=10):
    lo, hi = x.min(0), x.max(0)
    q = np.minimum(((x - lo) / (hi - lo) * (1 << bits)).astype(np.int64), (1 << bits) - 1)
    key = np.zeros(len(x), np.int64)
    for b in range(bits):
        for c in range(3):
            key |= ((q[:, c] >> b) & 1) << (3 * b + (2 - c))
    return np.argsort(key, kind="stable")


def kd_order(x, leaf):
    out = []

    def rec(ids):
        if len(ids) <= leaf:
            out.append(ids)
            return
        pts = x[ids]
        d = np.argmax(pts.max(0) - pts.min(0))
        h = (len(ids) // 2 + leaf - 1) // leaf * leaf
        if h >= len(ids):
            h = len(ids) // 2
        o = np.argpartition(pts[:, d], h - 1)
        rec(ids[o[:h]])
        rec(ids[o[h:]])

    rec(np.arange(len(x)))
    return np.concatenate(out)


def hybrid(m, mo, block, leaf):
    return np.concatenate([mo[s:s + block][kd_order(m[mo[s:s + block]], leaf)] for s in range(0, len(mo), block)])


def inside(q, r, lo, hi, chunk=1000):
    out = np.zeros(len(q), np.int64)
    for s in range(0, len(q), chunk):
        e = np.maximum(np.maximum(lo[None] - q[s:s + chunk, None], q[s:s + chunk, None] - hi[None]), 0)
        out[s:s + chunk] = ((e * e).sum(2) <= (r[s:s + chunk, None] * 1.000001) ** 2).sum(1)
    return out


def group_boxes(lo, hi, g):
    n = len(lo)
    k = (n + g - 1) // g
    L = np.full((k * g, 3), np.inf)
    H = np.full((k * g, 3), -np.inf)
    L[:n], H[:n] = lo, hi
    return L.reshape(k, g, 3).min(1), H.reshape(k, g, 3).max(1)


def wide_stats(m, order, q, r):
    x = m[order]
    lo, hi = group_boxes(x, x, 32)
    per_level = [inside(q, r, lo, hi).mean()]
    while len(lo) > 1:
        lo, hi = group_boxes(lo, hi, 32)
        per_level.append(inside(q, r, lo, hi).mean())
    internal = sum(per_level[2:])
    revisits = sum(per_level[1:-1])
    return per_level[0], per_level[1], per_level[1] + internal + revisits


def binary_stats(m, order, q, r):
    x = m[order]
    lo, hi = group_boxes(x, x, 8)
    tot = [inside(q, r, lo, hi, 300).mean()]
    while len(lo) > 1:
        lo, hi = group_boxes(lo, hi, 2)
        tot.append(inside(q, r, lo, hi, 300).mean())
    return tot[0], sum(tot[1:])


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "wide"
    rng = np.random.default_rng(0)
    if which == "wide":
        m = cloud("horse_ref")
        p = np.load(os.path.join(GOLDEN, "horse_tr1.npz"))["new_p"]
        sub = rng.choice(len(p), 4000, replace=False)
        q, r = p[sub], cKDTree(m).query(p[sub])[0]
        mo = morton_order(m)
        rows = [("morton", mo)] + [(f"median splits in {b}-point blocks", hybrid(m, mo, b, 32)) for b in (1024, 2048, 8192, 32768)]
        rows.append(("k-d from the top", kd_order(m, 32)))
        for name, order in rows:
            leaves, blocks, loads = wide_stats(m, order, q, r)
            print(f"{name:38s} leaves {leaves:5.2f}  level-0 blocks {blocks:4.2f}  block loads {loads:5.1f}")
    else:
        m, p, _ = synthetic_pair(1 << 17, seed=4)
        sub = rng.choice(len(p), 1500, replace=False)
        q, r = p[sub], cKDTree(m).query(p[sub])[0]
        mo = morton_order(m)
        rows = [("morton", mo), ("median splits in 2048-point blocks, down to 32", hybrid(m, mo, 2048, 32))]
        rows += [(f"median splits in {b}-point blocks", hybrid(m, mo, b, 8)) for b in (8192, 32768)]
        rows.append(("k-d from the top", kd_order(m, 8)))
        for name, order in rows:
            leaves, boxes = binary_stats(m, order, q, r)
            print(f"{name:48s} leaves {leaves:5.2f}  boxes over all levels {boxes:5.1f}")
