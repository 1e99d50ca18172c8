#!/bin/bash
# Round 2, first gpurun call (1 GPU): regression of the restored tree + every experiment staged at the end of round 1
# (ICP_GRAPH, ICP_NN_MIN_UNITS, ICP_KD_GLOBAL_MIN, the filter's build-time variants) + the evidence VERDICT r01 asked
# for: ncu --set full of nn_refine / nn_exact / align_fused (horse) and moments / transform / align_fused (2^22, tree
# mode), compute-sanitizer memcheck + racecheck.   BEFORE the call:  make -C iterative-closest-point_b200 <variants>
#   gpurun --timeout 2400 -- 'bash tools/r02_call1.sh'
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
timeout 600 python bench.py --workload horse --steps 20 --warmup 5 > $O/bench_horse.json 2> $O/bench_horse.err
timeout 600 python tools/graph_check.py > $O/graph_check.log 2>&1; echo "rc=$?" >> $O/graph_check.log
ICP_PROFILE=1 timeout 600 python tools/quick_bench.py > $O/quick_profile.log 2>&1
timeout 600 python tools/quick_bench.py > $O/quick.log 2>&1
# small clouds: fewer chunk-units per filter CTA than one staged tile (cow: 18 CTAs -> up to 144)
for u in 4 2 1; do
  ICP_NN_MIN_UNITS=$u timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/minunits_${u}_pytest.log 2>&1; echo "rc=$?" >> $O/minunits_${u}_pytest.log
  ICP_NN_MIN_UNITS=$u QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/minunits_${u}_quick.log 2>&1
done
QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/minunits_8_quick.log 2>&1
for u in 1 2; do   # with the graph on top
  ICP_GRAPH=1 ICP_NN_MIN_UNITS=$u QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/minunits_${u}_graph_quick.log 2>&1
done
# ---- ncu --set full of the kernels that had launch-list durations only ---------------------------
NCU="ncu --set full --clock-control none --import-source on"
timeout 600 $NCU -k "regex:nn_refine_kernel|nn_exact_kernel|align_fused_kernel" -s 6 -c 3 -o $O/full_horse_tail python tools/prof_step.py horse 4 > $O/ncu_horse_tail.log 2>&1
ICP_NN=tree ICP_ALIGN=split timeout 600 python tools/prof_step.py syn4m 2 > $O/prof_syn4m_tree_split.log 2>&1 && \
ICP_NN=tree ICP_ALIGN=split timeout 900 $NCU -k "regex:moments_kernel|transform_kernel" -s 2 -c 2 -o $O/full_syn4m_split python tools/prof_step.py syn4m 2 > $O/ncu_syn4m_split.log 2>&1
ICP_NN=tree ICP_ALIGN=fused timeout 900 $NCU -k "regex:align_fused_kernel" -s 1 -c 1 -o $O/full_syn4m_fused python tools/prof_step.py syn4m 2 > $O/ncu_syn4m_fused.log 2>&1
ICP_NN=tree timeout 300 python tools/prof_tail.py syn4m 3 > $O/prof_tail_syn4m.log 2>&1
timeout 300 python tools/prof_tail.py horse 30 > $O/prof_tail_horse.log 2>&1
# ---- filter geometry A/B ----------------------------------------------------------------------------
timeout 1500 bash tools/filter_ab.sh g128u8 g256u8 r12c3g128u4 r12c3g128u8 g128u4 u8 > $O/filter_ab.log 2>&1
# ---- tree search: global k-d levels also for models below 131 072 points -----------------------------
ICP_KD_GLOBAL_MIN=16384 timeout 900 python -m pytest tests/test_gpu_tree.py -m gpu -x -q > $O/kdmin_pytest.log 2>&1; echo "rc=$?" >> $O/kdmin_pytest.log
for w in horse horse_tr2 bunny syn64k; do
  timeout 300 python bench.py --nn tree --skip-cpu-baseline --workload $w --steps 10 --warmup 3 > $O/kdmin_default_$w.json 2>/dev/null
  ICP_KD_GLOBAL_MIN=16384 timeout 300 python bench.py --nn tree --skip-cpu-baseline --workload $w --steps 10 --warmup 3 > $O/kdmin_16k_$w.json 2>/dev/null
done
# ---- sanitizer ------------------------------------------------------------------------------------------
timeout 900 compute-sanitizer --tool memcheck python tools/sanitize_case.py > $O/sanitizer_memcheck.log 2>&1; echo "rc=$?" >> $O/sanitizer_memcheck.log
timeout 900 compute-sanitizer --tool racecheck python tools/sanitize_case.py > $O/sanitizer_racecheck.log 2>&1; echo "rc=$?" >> $O/sanitizer_racecheck.log
tail -3 $O/pytest_gpu.log; tail -1 $O/smoke.log; cat $O/graph_check.log; head -3 $O/minunits_*_quick.log; tail -2 $O/minunits_*_pytest.log $O/kdmin_pytest.log
grep -ho '"ms_per_step": [0-9.]*' $O/kdmin_*.json | paste - - ; cat $O/ab_summary.txt; tail -4 $O/sanitizer_*.log
