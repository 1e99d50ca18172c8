"""Quick device timings of the hot path (development aid; bench.py is the contract)."""
import os, sys, time, lzma
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "iterative-closest-point_b200")); sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import icp_b200
from conftest import synthetic_pair

def load(name):
    with lzma.open(os.path.join(ROOT, "tests/golden/clouds", name + ".txt.xz"), "rt") as f:
        lines = f.read().splitlines()[1:]
    return np.array([[float(x) for x in l.split(",")[:3]] for l in lines])

PEAK = 148 * 128 * 2 * 1.965e9
def closest_timing(tag, m, p, reps=5):
    ctx = icp_b200.Context(0); ctx.set_model(m); ctx.set_scene(p)
    ctx.closest(want_idx=False, want_Y=False)
    best = 1e9
    for _ in range(reps):
        ctx.closest(want_idx=False, want_Y=False)
        best = min(best, ctx.stats()["ms_total"])
    st = ctx.stats()
    ev = len(m) * len(p) / (best * 1e-3)
    print(f"{tag}: closest {best:.3f} ms  {ev/1e12:.3f} T evals/s  {ev*8/PEAK*100:.1f}% of FP32 roofline  splits={st['n_splits']} "
          f"rerank_full/pass={st['rerank_full']/st['nn_calls']:.1f} of {len(p)}", flush=True)
    ctx.close()

def loop_timing(tag, m, p, max_iter=30, check=True):
    ctx = icp_b200.Context(0); ctx.set_model(m); ctx.set_scene(p)
    ctx.run(max_iter, check_sizes=check); 
    for _ in range(3):
        ctx.reset_scene(); t0 = time.perf_counter(); it, hist = ctx.run(max_iter, check_sizes=check); dt = time.perf_counter() - t0
    st = ctx.stats()
    print(f"{tag}: loop {it} it  dev {st['ms_total']:.3f} ms (filter {st['ms_filter']:.3f} refine {st['ms_refine']:.3f} exact {st['ms_exact']:.3f} align {st['ms_align']:.3f} ms) wall {dt*1e3:.3f} ms  "
          f"{it/(st['ms_total']*1e-3):.1f} it/s  {it*len(m)*len(p)/(st['ms_total']*1e-3)/1e12:.3f} T evals/s  last err {hist[-1]:.3e}", flush=True)
    if os.environ.get("ICP_PROFILE"):
        if tag == "cow":
            bt = ctx.block_trace()
            print("   per block (SM, entry, barrier, solve, exit us):", " | ".join(" ".join(f"{v:.1f}" for v in r) for r in bt if r[4] >= 0))
        print("   align trace (us: gap after K1 | from entry: moments, partials, barrier, reduce, solve, transform, last block out | partials re-read, block sum):", np.round(ctx.trace(), 2))
    ctx.close()

if __name__ == "__main__":
    cow_r, cow_1 = load("cow_ref"), load("cow_tr1")
    closest_timing("cow", cow_r, cow_1); loop_timing("cow", cow_r, cow_1)
    if os.environ.get("QB_ONLY_SMALL"):          # small clouds only
        for n in (8192, 16384):
            m, p, _ = synthetic_pair(n, seed=n)
            closest_timing(f"synthetic {n}", m, p); loop_timing(f"synthetic {n}", m, p, max_iter=10)
        sys.exit(0)
    h_r, h_1 = load("horse_ref"), load("horse_tr1")
    closest_timing("horse", h_r, h_1); loop_timing("horse_tr1", h_r, h_1)
    for lg in (18, 20):
        m, p, _ = synthetic_pair(1 << lg, seed=lg)
        closest_timing(f"synthetic 2^{lg}", m, p, reps=3)
