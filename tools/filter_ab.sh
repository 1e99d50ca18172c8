#!/bin/bash
# One gpurun call: A/B of the filter kernel's compile-time geometry (make variants; icp_common.cuh ICP_NN_*).
# Every variant first has to pass the index-parity tests (the chunk size and the tile size enter nn_refine and the
# stream-K arithmetic), then gets bench lines on horse (one wave, 61 chunks per CTA) and on 2^18 points (long ranges).
# The default build runs first AND last so that drift of the box shows.   usage: bash tools/filter_ab.sh [variants...]
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
V="${@:-r8c4g64u2 r8c4g128u8 r12c3g256u8 r12c3g128u16 r16c2g128u8 r10c3g128u8 rf4 rf16 inp1 inp4}"
LIBDIR=$PWD/iterative-closest-point_b200/lib
run() {   # name, library
  export ICP_B200_LIB=$2
  timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/ab_$1_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/ab_$1_pytest.log
  for w in horse syn256k; do
    timeout 300 python bench.py --skip-cpu-baseline --workload $w --steps 20 --warmup 3 > gpurun_out/ab_$1_$w.json 2> gpurun_out/ab_$1_$w.err
  done
  # K2 / K4 as HBM-bound passes on 2^22 points (tree search so that the run takes milliseconds): alignment variants only
  case $1 in fused*|default_a) ICP_NN=tree timeout 300 python tools/prof_tail.py syn4m 3 > gpurun_out/ab_$1_tail.log 2>&1;; esac
}
run default_a $LIBDIR/libicp_b200.so
for v in $V; do [ -f $LIBDIR/variants/libicp_b200_$v.so ] && run $v $LIBDIR/variants/libicp_b200_$v.so; done
run default_b $LIBDIR/libicp_b200.so
python - $V <<'PY' | tee gpurun_out/ab_summary.txt
import json, sys
print("%-14s %-8s %10s %10s %8s  %s" % ("variant", "workload", "ms/step", "T evals/s", "K1 frac", "parity"))
for v in ["default_a"] + sys.argv[1:] + ["default_b"]:
    try:
        par = open("gpurun_out/ab_%s_pytest.log" % v).read().strip().splitlines()[-2:]
    except OSError:
        continue
    for w in ("horse", "syn256k"):
        try:
            d = json.loads(open("gpurun_out/ab_%s_%s.json" % (v, w)).read().strip().splitlines()[-1])
            print("%-14s %-8s %10.3f %10.3f %8.3f  %s" % (v, w, d["ms_per_step"], d["value"] / 1e12, d["roofline"]["frac"], " | ".join(par)))
        except Exception as e:
            print("%-14s %-8s failed: %s" % (v, w, e))
PY
grep -H "K2+K4" gpurun_out/ab_*_tail.log | tee -a gpurun_out/ab_summary.txt
