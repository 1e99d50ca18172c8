"""ICP_GRAPH=1 (icp_run re-launching one captured iteration as a CUDA graph) against the plain enqueue path, on one GPU:
identical iteration counts, error histories and new_p bit for bit on cow / horse_tr2 / a synthetic cloud, then loop times.
Experiment of DESIGN.md §9 (launch gaps); not part of the test suite until it has run on a GPU once."""
import os, sys, lzma, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "iterative-closest-point_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import synthetic_pair


def load(name):
    with lzma.open(os.path.join(ROOT, "tests/golden/clouds", name + ".txt.xz"), "rt") as f:
        lines = f.read().splitlines()[1:]
    return np.array([[float(x) for x in l.split(",")[:3]] for l in lines])


def run(mode, m, p, max_iter, reps=5):
    os.environ["ICP_GRAPH"] = str(mode)
    import icp_b200
    ctx = icp_b200.Context(0)           # ICP_GRAPH is read at icp_create
    ctx.set_model(m); ctx.set_scene(p)
    it, hist = ctx.run(max_iter)
    newp = ctx.get_scene()
    best = 1e9
    for _ in range(reps):
        ctx.reset_scene()
        it2, hist2 = ctx.run(max_iter)
        assert it2 == it and np.array_equal(hist, hist2)
        best = min(best, ctx.stats()["ms_total"])
    launches = ctx.stats()["kernels_launched"]
    ctx.close()
    return it, np.array(hist), newp, best, launches


if __name__ == "__main__":
    cases = [("cow", load("cow_ref"), load("cow_tr1"), 30), ("horse_tr2", load("horse_ref"), load("horse_tr2"), 30),
             ("horse_tr1", load("horse_ref"), load("horse_tr1"), 30)]
    m, p, _ = synthetic_pair(1 << 16, seed=16)
    cases.append(("syn64k", m, p, 10))
    for name, m, p, mi in cases:
        a = run(0, m, p, mi)
        b = run(1, m, p, mi)
        same = a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
        print(f"{name}: iters {a[0]} / {b[0]}  identical={same}  plain {a[3]:.3f} ms ({a[4]} launches)  graph {b[3]:.3f} ms ({b[4]} launches)", flush=True)
        assert same, name
