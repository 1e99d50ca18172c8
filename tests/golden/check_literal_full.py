"""One-off check, run in the build container (70 s on 8 threads): the oracle's LITERAL nearest-neighbour mode
(sqrt(pow(dx,2)+pow(dy,2)+pow(dz,2)) with libm's pow really called, as the reference's Eigen expression does) against
the x*x form the CUDA path evaluates, on the FULL first pass of every large reference cloud pair.
Result of 2026-11-05 (glibc 2.39, pow(x,2) != x*x for 0.090 % of inputs): 0 index mismatches on
cow_tr2 (2 903^2 pairs), horse_tr1 and horse_tr2 (48 485^2), bun045 x bun000 (40 097 x 40 256) -> literal_full_check.json."""
import json, os
import sys, lzma, time, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import oracle as O
def load(name):
    with lzma.open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "clouds", name + ".txt.xz"), "rt") as f:
        lines = f.read().splitlines()[1:]
    return np.array([[float(x) for x in l.split(",")[:3]] for l in lines])
out = {}
for mn, pn in [("cow_ref","cow_tr2"),("horse_ref","horse_tr1"),("horse_ref","horse_tr2"),("bun000","bun045")]:
    m, p = load(mn), load(pn)
    t0=time.time(); i_lit,_ = O.closest_matrix(p, m, nn_mode=O.NN_LITERAL); t1=time.time()
    i_mul,_ = O.closest_matrix(p, m, nn_mode=O.NN_MULSQRT); t2=time.time()
    out[f"{pn} x {mn}"] = {"scene_points": len(p), "model_points": len(m), "index_mismatches": int((i_lit != i_mul).sum())}
    print(f"{pn} x {mn}: {len(p)} x {len(m)} pairs, literal {t1-t0:.1f}s mulsqrt {t2-t1:.1f}s, index mismatches {int((i_lit!=i_mul).sum())}", flush=True)
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "literal_full_check.json"), "w"), indent=1)
