#!/bin/bash
# Round 2, third gpurun call (1 GPU): slot-pruned rescans of undecided points (in nn_refine / nn_exact), the wide tree
# kernel's warm-start fix, per-block trace of the fused alignment kernel on cow, more variants, config 5 at 1 024 pairs.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
ICP_PROFILE=1 timeout 600 python tools/quick_bench.py > $O/quick_profile.log 2>&1
timeout 600 python tools/quick_bench.py > $O/quick.log 2>&1
timeout 300 python tools/prof_tail.py horse 30 > $O/prof_tail_horse.log 2>&1
timeout 300 python tools/prof_tail.py syn256k 5 > $O/prof_tail_syn256k.log 2>&1
timeout 300 python tools/prof_tail.py syn1m 2 > $O/prof_tail_syn1m.log 2>&1
ICP_NN=tree timeout 300 python tools/prof_tail.py syn4m 3 > $O/prof_tail_syn4m.log 2>&1
timeout 600 python bench.py --workload horse --steps 20 --warmup 5 > $O/bench_horse.json 2> $O/bench_horse.err
timeout 600 python bench.py --workload cow --steps 20 --warmup 5 --skip-cpu-baseline > $O/bench_cow.json 2> $O/bench_cow.err
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_default.json 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
bash tools/checked_run.sh > $O/checked_run.log 2>&1
timeout 1500 bash tools/filter_ab.sh r16c2g128u8 r16c2g128u16 r16c2g256u8 r12c3g128u16 fused3 fused4 rf4 rf16 > $O/filter_ab.log 2>&1
timeout 900 python bench.py --workload batch64k --pairs 1024 --steps 2 --warmup 1 > $O/batch64k_1024_n1.json 2> $O/batch64k_1024_n1.err
tail -n 3 $O/pytest_gpu.log; tail -n 1 $O/smoke.log; head -12 $O/quick_profile.log; cat $O/prof_tail_*.log; cat $O/checked_run.log; cat $O/ab_summary.txt
cut -c1-600 $O/bench_default.json $O/bench_horse.json $O/bench_cow.json $O/batch64k_1024_n1.json
