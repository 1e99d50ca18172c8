#!/usr/bin/env python3
"""Static statistics of nn_filter_kernel in one or more builds of libicp_b200.so (no GPU needed).

For each library: registers / stack, the inner (q) loop's instruction mix, and the instructions one thread executes
per 64 model points x NN_R scene points (= q-loop trips + the chunk-tracking block between two trips of the c loop),
i.e. the share of issue slots that the formulation's FFMA2 / FMNMX3 can reach at best.  DESIGN.md §4 (K1) argues
that every warp instruction of this loop costs the same two issue cycles, so this share IS the kernel's ceiling as a
fraction of the 8-flop roofline; the measured figure additionally loses staging, prologue, tails and padding.

    python tools/sass_filter_stats.py iterative-closest-point_b200/lib/libicp_b200.so iterative-closest-point_b200/lib/variants/*.so
"""
import re
import subprocess
import sys

INSTR = re.compile(r"^\s+/\*([0-9a-f]{4,})\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)(.*?);")


def kernel_sass(lib, name="nn_filter_kernel"):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    cur, res = False, []
    for line in out.splitlines():
        if "Function :" in line:
            cur = name in line
            continue
        m = INSTR.match(line)
        if cur and m:
            res.append((int(m.group(1), 16), m.group(2), m.group(3)))
    return res


def resources(lib, name="nn_filter_kernel"):
    out = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True, check=True).stdout
    hit = False
    for line in out.splitlines():
        if "Function " in line:
            hit = name in line
        r = re.search(r"REG:(\d+) STACK:(\d+)", line)
        if r and hit:
            return int(r.group(1)), int(r.group(2))
    return None, None


def loops(instrs):
    """All backward branches as (start index, end index), innermost first."""
    addr = {a: i for i, (a, _, _) in enumerate(instrs)}
    res = []
    for i, (a, op, args) in enumerate(instrs):
        t = re.search(r"0x([0-9a-f]+)", args) if op.startswith("BRA") else None
        if t and int(t.group(1), 16) < a and int(t.group(1), 16) in addr:
            res.append((addr[int(t.group(1), 16)], i))
    return sorted(res, key=lambda r: r[1] - r[0])


def cnt(body, prefix):
    return sum(1 for _, op, _ in body if op.startswith(prefix))


def stats(lib):
    k = kernel_sass(lib)
    reg, stack = resources(lib)
    ls = [l for l in loops(k) if cnt(k[l[0]:l[1] + 1], "FFMA2")]
    q0, q1 = ls[0]
    qbody = k[q0:q1 + 1]
    # the c loop: the smallest loop that strictly contains the q loop (None when the q loop is fully unrolled away)
    outer = next(((a, b) for a, b in ls[1:] if a <= q0 and b >= q1 and (a, b) != (q0, q1)), None)
    lds, f2, m3 = cnt(qbody, "LDS.128"), cnt(qbody, "FFMA2"), cnt(qbody, "FMNMX3")
    spills = cnt(qbody, "LDL") + cnt(qbody, "STL")
    r_points = f2 // (3 * lds // 2) if lds else 0      # scene points per thread: FFMA2 per LDS quadruple / 6
    d = dict(lib=lib.split("/")[-1], reg=reg, stack=stack, q_len=len(qbody), lds=lds, ffma2=f2, fmnmx3=m3,
             q_other=len(qbody) - lds - f2 - m3, q_spills=spills, r=r_points)
    if outer:
        cbody = k[outer[0]:outer[1] + 1]
        quads_per_trip = lds // 4
        # chunk size is not visible in the SASS; report per-trip and per-c-iteration overhead separately
        d["c_overhead"] = len(cbody) - len(qbody)
        d["c_spills"] = cnt(cbody, "LDL") + cnt(cbody, "STL") - spills
        d["quads_per_trip"] = quads_per_trip
    return d


def main():
    libs = sys.argv[1:]
    if not libs:
        print(__doc__)
        return 1
    print(f"{'library':28s} {'G':>4s} {'reg':>4s} {'stk':>4s} {'R':>3s} {'q-loop':>7s} {'LDS':>4s} {'FFMA2':>6s} {'FMNMX3':>7s} "
          f"{'other':>6s} {'spill':>6s} {'c-ovh':>6s} {'formulation share':>18s}")
    for lib in libs:
        d = stats(lib)
        form_trip = d["ffma2"] + d["fmnmx3"]
        g = re.search(r"g(\d+)", d["lib"].replace("libicp_b200", ""))      # chunk size: from the variant's name (default 64)
        G = int(g.group(1)) if g else 64
        trips = (G // 4) / max(d.get("quads_per_trip", 1), 1)
        total = trips * d["q_len"] + d.get("c_overhead", 0)
        share = trips * form_trip / total if total else float("nan")
        print(f"{d['lib']:28s} {G:4d} {d['reg']:4d} {d['stack']:4d} {d['r']:3d} {d['q_len']:7d} {d['lds']:4d} {d['ffma2']:6d} "
              f"{d['fmnmx3']:7d} {d['q_other']:6d} {d['q_spills']:6d} {d.get('c_overhead', -1):6d} {share:18.3f}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
