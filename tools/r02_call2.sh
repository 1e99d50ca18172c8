#!/bin/bash
# Round 2, second gpurun call (1 GPU): the new default build (filter geometry r12c3g128u8, undecided points rescanned
# inside nn_refine, 32-byte model records for every gather, two points in flight in K2/K4), the new bench default
# (2^22 synthetic, secondary.horse), more filter / refine variants, the checked build, ncu evidence.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_default.json 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
timeout 600 python bench.py --workload horse --steps 20 --warmup 5 > $O/bench_horse.json 2> $O/bench_horse.err
timeout 600 python bench.py --workload cow --steps 20 --warmup 5 --skip-cpu-baseline > $O/bench_cow.json 2> $O/bench_cow.err
ICP_PROFILE=1 timeout 600 python tools/quick_bench.py > $O/quick_profile.log 2>&1
timeout 600 python tools/quick_bench.py > $O/quick.log 2>&1
ICP_NN=tree timeout 300 python tools/prof_tail.py syn4m 3 > $O/prof_tail_syn4m.log 2>&1
timeout 300 python tools/prof_tail.py horse 30 > $O/prof_tail_horse.log 2>&1
timeout 300 python tools/prof_tail.py syn256k 5 > $O/prof_tail_syn256k.log 2>&1
bash tools/checked_run.sh > $O/checked_run.log 2>&1
# ---- ncu -----------------------------------------------------------------------------------------------
NCU="ncu --set full --clock-control none --import-source on"
ICP_NN=tree ICP_ALIGN=split timeout 900 $NCU -k "regex:moments_kernel|transform_kernel" -s 2 -c 2 -o $O/full_syn4m_split python tools/prof_step.py syn4m 2 > $O/ncu_syn4m_split.log 2>&1
timeout 600 $NCU -k "regex:nn_filter_kernel|nn_refine_kernel|align_fused_kernel" -s 9 -c 3 -o $O/full_horse python tools/prof_step.py horse 4 > $O/ncu_horse.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_horse.csv python tools/prof_step.py horse 3 > $O/ncu_launch.log 2>&1
timeout 900 $NCU -k "regex:nn_filter_kernel" -s 1 -c 1 -o $O/full_syn4m_filter python tools/prof_step.py syn4m 1 > $O/ncu_syn4m_filter.log 2>&1
# ---- A/B ------------------------------------------------------------------------------------------------
timeout 1500 bash tools/filter_ab.sh > $O/filter_ab.log 2>&1
tail -n 3 $O/pytest_gpu.log; tail -n 1 $O/smoke.log; cat $O/bench_default.json | cut -c1-1500; cat $O/quick_profile.log | head -8; cat $O/prof_tail_*.log; cat $O/checked_run.log; cat $O/ab_summary.txt
