"""One-off check, run in the build container (~25 min on 8 threads): WHOLE ICP loops in the oracle's LITERAL modes
(libm pow really called, the reference's max_element_index rule) against the committed golden vectors, which were
generated in the mulsqrt / intended modes the CUDA path implements: correspondence indices of EVERY iteration (CRC),
iteration counts, error values and the final new_p.  Writes literal_loops_check.json."""
import json, lzma, os, sys, time, zlib
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle as O


def load(name):
    with lzma.open(os.path.join(HERE, "clouds", name + ".txt.xz"), "rt") as f:
        lines = f.read().splitlines()[1:]
    return np.array([[float(x) for x in l.split(",")[:3]] for l in lines])


meta = json.load(open(os.path.join(HERE, "meta.json")))["cases"]
out = {}
for case in sys.argv[1:] or ["cow_tr1", "cow_tr2", "horse_tr2", "bunny", "horse_tr1"]:
    c = meta[case]
    g = np.load(os.path.join(HERE, case + ".npz"))
    m, p = load(c["model"]), load(c["scene"])
    t0 = time.time()
    r = O.icp_run(m, p, c["max_iter"], nn_mode=O.NN_LITERAL, eig_mode=O.EIG_LITERAL, allow_unequal=c["allow_unequal"], keep_idx=True)
    crc = [zlib.crc32(r["idx"][i].tobytes()) for i in range(r["iters"])]
    same_idx = r["iters"] == int(g["iters"]) and all(int(a) == int(b) for a, b in zip(crc, g["idx_crc"]))
    rel_err = float(np.max(np.abs(r["err"] - g["err"]) / g["err"])) if r["iters"] == int(g["iters"]) else None
    dp = float(np.max(np.abs(r["new_p"] - g["new_p"]))) if r["iters"] == int(g["iters"]) else None
    out[case] = {"iterations": int(r["iters"]), "golden_iterations": int(g["iters"]), "indices_identical_every_iteration": bool(same_idx),
                 "max_rel_diff_error_value": rel_err, "max_abs_diff_new_p": dp, "seconds": round(time.time() - t0, 1)}
    print(case, out[case], flush=True)
    json.dump(out, open(os.path.join(HERE, "literal_loops_check.json"), "w"), indent=1)
