"""Generates the committed fixtures under tests/golden/ from the reference's data and the CPU oracle.

Run in the build container only (it reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_golden.py

Outputs
  clouds/<name>.txt.xz   the reference's point clouds (data_students/*.txt), xz-compressed verbatim, so
                         that the loader is exercised on the real format (header line, 5-column cow_tr2)
  <case>.npz             oracle results of the full loop (src/cpu.cc:55-79): iteration count, the printed
                         error values, s/R/t per iteration, correspondence indices of iteration 0 and
                         the last iteration, CRC32 of the indices of every iteration, final new_p
  meta.json              oracle modes used, literal-vs-mulsqrt agreement, eigen-rule agreement
"""
import json
import lzma
import os
import sys
import time
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

REF_DATA = "/root/reference/data_students"
CLOUDS = ["cow_ref", "cow_tr1", "cow_tr2", "horse_ref", "horse_tr1", "horse_tr2", "bun000", "bun045"]
CASES = {  # case -> (model, scene, max_iter, allow_unequal)
    "cow_tr1": ("cow_ref", "cow_tr1", 30, False),
    "cow_tr2": ("cow_ref", "cow_tr2", 30, False),
    "horse_tr1": ("horse_ref", "horse_tr1", 30, False),
    "horse_tr2": ("horse_ref", "horse_tr2", 30, False),
    "bunny": ("bun000", "bun045", 30, True),
}


def main():
    os.makedirs(os.path.join(HERE, "clouds"), exist_ok=True)
    pts = {}
    for name in CLOUDS:
        src = os.path.join(REF_DATA, name + ".txt")
        raw = open(src, "rb").read()
        with lzma.open(os.path.join(HERE, "clouds", name + ".txt.xz"), "wb", preset=9 | lzma.PRESET_EXTREME) as f:
            f.write(raw)
        pts[name] = O.load_matrix(src)
        print(name, pts[name].shape)
    meta = {"nn_mode": "mulsqrt", "eig_mode": "intended", "threads": O.num_threads(), "cases": {}}
    only = sys.argv[1:] or list(CASES)
    for case in only:
        mname, pname, max_iter, allow = CASES[case]
        m, p = pts[mname], pts[pname]
        t0 = time.time()
        r = O.icp_run(m, p, max_iter, nn_mode=O.NN_MULSQRT, eig_mode=O.EIG_INTENDED, allow_unequal=allow, keep_idx=True)
        dt = time.time() - t0
        crc = np.array([zlib.crc32(r["idx"][i].tobytes()) for i in range(r["iters"])], np.uint32)
        np.savez_compressed(os.path.join(HERE, case + ".npz"), iters=r["iters"], err=r["err"], s=r["s"], R=r["R"],
                            t=r["t"], idx_first=r["idx"][0], idx_last=r["idx"][-1], idx_crc=crc,
                            new_p=r["new_p"])
        info = {"model": mname, "scene": pname, "max_iter": max_iter, "allow_unequal": allow,
                "iters": int(r["iters"]), "err_last": float(r["err"][-1]), "oracle_seconds": round(dt, 1)}
        # literal (libm pow + sqrt, src/cpu.cc:17-19) against the x*x form on the first pass
        sub = slice(0, min(len(p), 4096))
        il, _ = O.closest_matrix(p[sub], m, O.NN_LITERAL)
        info["literal_first_pass_mismatches"] = int((il != r["idx"][0][sub]).sum())
        info["literal_first_pass_points"] = int(len(il))
        # literal eigen rule (general solver + max_element_index bug) against the intended one
        rl = O.icp_run(m, p, min(max_iter, 3), nn_mode=O.NN_MUL, eig_mode=O.EIG_LITERAL, allow_unequal=allow)
        info["literal_eig_err_first3"] = [float(x) for x in rl["err"]]
        info["intended_err_first3"] = [float(x) for x in r["err"][:3]]
        meta["cases"][case] = info
        print(case, info)
    mp = os.path.join(HERE, "meta.json")
    if os.path.exists(mp) and sys.argv[1:]:
        old = json.load(open(mp))
        old["cases"].update(meta["cases"])
        meta = old
    json.dump(meta, open(mp, "w"), indent=1)


if __name__ == "__main__":
    main()
