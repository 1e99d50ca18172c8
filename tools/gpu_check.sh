#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench (both arms), quick timings, ncu launch list (+ optional full capture).
# usage: bash tools/gpu_check.sh [full-capture kernel regex]
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_horse.json 2> gpurun_out/bench_horse.err; echo "rc=$?" >> gpurun_out/bench_horse.err
timeout 600 python tools/quick_bench.py > gpurun_out/quick.log 2>&1
timeout 600 python tools/prof_step.py horse 3 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_horse.csv python tools/prof_step.py horse 3 > gpurun_out/ncu_launch.log 2>&1
if [ -n "$1" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$1" -s 2 -c 1 -o gpurun_out/full_$1 python tools/prof_step.py horse 3 > gpurun_out/ncu_full.log 2>&1
fi
tail -3 gpurun_out/pytest_gpu.log
