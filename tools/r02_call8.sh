#!/bin/bash
# Round 2, 1-GPU call: second version of the one-launch loop kernel (4 scene points per warp pass, one grid barrier per
# iteration): its own tests first under a short timeout, then the whole GPU suite, timings against the per-kernel loop.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "small_loop or full_loop_matches_golden or non_finite" > $O/pytest_small.log 2>&1; echo "rc=$?" >> $O/pytest_small.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/quick_small.log 2>&1
ICP_SMALL=0 QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/quick_small_off.log 2>&1
timeout 300 python bench.py --workload cow --steps 20 --warmup 5 > $O/wl_cow.json 2> $O/wl_cow.err
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
for f in cow_ref cow_tr1; do mkdir -p /tmp/cl; xz -dc tests/golden/clouds/$f.txt.xz > /tmp/cl/$f.txt; done
iterative-closest-point_b200/bin/bench /tmp/cl/cow_ref.txt /tmp/cl/cow_tr1.txt > $O/cppbench_cow.txt 2>/dev/null
bash tools/gpu_refsrc_check.sh > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:icp_small_loop_kernel" -s 1 -c 1 -o $O/full_cow_small python tools/prof_step.py cow > $O/ncu_cow_small.log 2>&1
tail -n 5 $O/pytest_small.log; tail -n 1 $O/smoke.log; cat $O/quick_small.log $O/quick_small_off.log; cut -c1-250 $O/wl_cow.json; tail -n 6 $O/pytest_gpu.log; grep -E "opti_gpu_loop|resident_loop" $O/cppbench_cow.txt; grep -E "loop|IDENT" $O/refsrc_check.txt
