#!/bin/bash
# Round 2, 1-GPU call: the search kernels record the matched model point (SceneDev::y64*), K2 / K4 stream it instead of
# gathering m64[idx].  Whole suite + checked build, K2+K4 times and DRAM traffic on 2^22 points, no regression elsewhere.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
bash tools/checked_run.sh > $O/checked_run.log 2>&1
ICP_NN=tree timeout 300 python tools/prof_tail.py syn4m 3 > $O/prof_tail_syn4m.log 2>&1
ICP_NN=brute timeout 300 python tools/prof_tail.py horse 30 > $O/prof_tail_horse.log 2>&1
ICP_NN=brute timeout 300 python tools/prof_tail.py syn256k 5 > $O/prof_tail_syn256k.log 2>&1
ICP_NN=brute timeout 300 python tools/prof_tail.py syn1m 2 > $O/prof_tail_syn1m.log 2>&1
timeout 600 python bench.py --workload horse --steps 20 --warmup 5 --skip-cpu-baseline > $O/wl_horse.json 2> $O/wl_horse.err
timeout 600 python bench.py --nn tree --steps 10 --warmup 3 --skip-cpu-baseline --no-secondary > $O/tree_wl_syn4m.json 2> $O/tree_wl_syn4m.err
timeout 600 python bench.py --nn tree --workload horse --steps 10 --warmup 3 --skip-cpu-baseline > $O/tree_wl_horse.json 2> $O/tree_wl_horse.err
timeout 600 python bench.py --steps 3 --warmup 3 --skip-cpu-baseline > $O/bench_default.json 2> $O/bench_default.err
NCU="ncu --set full --clock-control none --import-source on"
ICP_NN=tree ICP_ALIGN=split timeout 900 $NCU -k "regex:moments_kernel|transform_kernel" -s 2 -c 2 -o $O/full_syn4m_split python tools/prof_step.py syn4m 2 > $O/ncu_syn4m_split.log 2>&1
ICP_NN=tree timeout 900 $NCU -k "regex:align_fused_kernel" -s 1 -c 1 -o $O/full_syn4m_fused python tools/prof_step.py syn4m 2 > $O/ncu_syn4m_fused.log 2>&1
ICP_NN=brute timeout 600 $NCU -k "regex:nn_refine_kernel|align_fused_kernel" -s 6 -c 2 -o $O/full_horse_tail python tools/prof_step.py horse 4 > $O/ncu_horse_tail.log 2>&1
tail -n 4 $O/pytest_gpu.log; tail -n 1 $O/smoke.log; cat $O/checked_run.log | tail -8; cat $O/prof_tail_*.log
for f in wl_horse tree_wl_syn4m tree_wl_horse bench_default; do cut -c1-230 $O/$f.json; done
