#!/bin/bash
# Round 2, 1-GPU call: the one-launch loop kernel for small clouds (tests, timings against the per-kernel loop), the final
# single-GPU numbers of every BASELINE config (both search modes), the C++ harnesses, and ncu evidence for every kernel.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out /tmp/cl
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/quick_small.log 2>&1
ICP_SMALL=0 QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/quick_small_off.log 2>&1
timeout 600 python tools/quick_bench.py > $O/quick.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_default.json 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
for w in cow horse horse_tr2 bunny syn64k syn256k syn1m; do
  st=10; [ $w = syn1m ] && st=3
  timeout 600 python bench.py --workload $w --steps $st --warmup 3 > $O/wl_$w.json 2> $O/wl_$w.err
  timeout 600 python bench.py --nn tree --workload $w --steps $st --warmup 3 --skip-cpu-baseline > $O/tree_wl_$w.json 2> $O/tree_wl_$w.err
done
ICP_SMALL=0 timeout 300 python bench.py --workload cow --steps 10 --warmup 3 --skip-cpu-baseline > $O/wl_cow_kernels.json 2> $O/wl_cow_kernels.err
timeout 600 python bench.py --nn tree --steps 10 --warmup 3 --skip-cpu-baseline --no-secondary > $O/tree_wl_syn4m.json 2> $O/tree_wl_syn4m.err
timeout 300 python tools/prof_tail.py horse 30 > $O/prof_tail_horse.log 2>&1
for f in cow_ref cow_tr1 horse_ref horse_tr1; do xz -dc tests/golden/clouds/$f.txt.xz > /tmp/cl/$f.txt; done
iterative-closest-point_b200/bin/bench /tmp/cl/cow_ref.txt /tmp/cl/cow_tr1.txt > $O/cppbench_cow.txt 2>/dev/null
iterative-closest-point_b200/bin/bench /tmp/cl/horse_ref.txt /tmp/cl/horse_tr1.txt > $O/cppbench_horse.txt 2>/dev/null
bash tools/gpu_refsrc_check.sh > /dev/null 2>&1
bash tools/checked_run.sh > $O/checked_run.log 2>&1
# ---- ncu: --set full of every kernel of the path, launch list of the horse loop -----------------------------
NCU="ncu --set full --clock-control none --import-source on"
timeout 600 $NCU -k "regex:nn_filter_kernel|nn_refine_kernel|align_fused_kernel" -s 9 -c 3 -o $O/full_horse python tools/prof_step.py horse 4 > $O/ncu_horse.log 2>&1
timeout 600 $NCU -k "regex:nn_exact_kernel" -s 1 -c 1 -o $O/full_syn256k_exact python tools/prof_step.py syn256k 2 > $O/ncu_syn256k.log 2>&1
ICP_NN=tree ICP_ALIGN=split timeout 900 $NCU -k "regex:moments_kernel|transform_kernel" -s 2 -c 2 -o $O/full_syn4m_split python tools/prof_step.py syn4m 2 > $O/ncu_syn4m_split.log 2>&1
ICP_NN=tree timeout 900 $NCU -k "regex:align_fused_kernel" -s 1 -c 1 -o $O/full_syn4m_fused python tools/prof_step.py syn4m 2 > $O/ncu_syn4m_fused.log 2>&1
timeout 600 $NCU -k "regex:icp_small_loop_kernel" -s 1 -c 1 -o $O/full_cow_small python tools/prof_step.py cow > $O/ncu_cow_small.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_horse.csv python tools/prof_step.py horse 3 > $O/ncu_launch.log 2>&1
timeout 900 $NCU -k "regex:nn_filter_kernel" -s 1 -c 1 -o $O/full_syn4m_filter python tools/prof_step.py syn4m 1 > $O/ncu_syn4m_filter.log 2>&1
tail -n 4 $O/pytest_gpu.log; tail -n 1 $O/smoke.log; cat $O/quick_small.log $O/quick_small_off.log; cut -c1-300 $O/bench_default.json; cat $O/cppbench_cow.txt | tail -16; cat $O/refsrc_check.txt | tail -22; cat $O/checked_run.log
for f in $O/wl_*.json $O/tree_wl_*.json; do python - <<PY
import json
try:
    d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f'.split('/')[-1], round(d['value']/1e12,3), 'T', round(d['ms_per_step'],3), 'ms e2e', round(d['e2e']['ms_per_step'],3), 'it', d['iterations_per_step'], 'frac', (d.get('roofline') or {}).get('frac'), 'cpu', (d.get('cpu_baseline') or {}).get('value'))
except Exception as e: print('$f FAILED', e)
PY
done
