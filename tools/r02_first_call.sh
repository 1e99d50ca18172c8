#!/bin/bash
# First gpurun call of the next round: regression (GPU tests, smoke, default bench) and the two experiments staged at
# the end of round 1 without GPU time: ICP_GRAPH (tools/graph_check.py) and the filter's build-time variants
# (tools/filter_ab.sh).  BEFORE the call, in the build container:  make -C iterative-closest-point_b200 variants
#   gpurun --timeout 3000 -- 'bash tools/r02_first_call.sh'
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_horse.json 2> gpurun_out/bench_horse.err
timeout 600 python tools/graph_check.py > gpurun_out/graph_check.log 2>&1; echo "rc=$?" >> gpurun_out/graph_check.log
ICP_PROFILE=1 timeout 600 python tools/quick_bench.py > gpurun_out/quick_profile.log 2>&1
timeout 2400 bash tools/filter_ab.sh > gpurun_out/filter_ab.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log; cat gpurun_out/graph_check.log; cat gpurun_out/ab_summary.txt
