#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench (both arms), quick timings, the C++ capture list, ncu launch list
# (+ optional full capture of one kernel).   usage: bash tools/gpu_check.sh [full-capture kernel regex]
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out /tmp/cl
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_horse.json 2> gpurun_out/bench_horse.err; echo "rc=$?" >> gpurun_out/bench_horse.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
timeout 600 python tools/quick_bench.py > gpurun_out/quick.log 2>&1
ICP_PROFILE=1 timeout 600 python tools/quick_bench.py > gpurun_out/quick_profile.log 2>&1
for f in cow_ref cow_tr1 horse_ref horse_tr1; do xz -dc tests/golden/clouds/$f.txt.xz > /tmp/cl/$f.txt; done
iterative-closest-point_b200/bin/bench /tmp/cl/cow_ref.txt /tmp/cl/cow_tr1.txt > gpurun_out/cppbench_cow.txt 2>/dev/null
iterative-closest-point_b200/bin/bench /tmp/cl/horse_ref.txt /tmp/cl/horse_tr1.txt > gpurun_out/cppbench_horse.txt 2>/dev/null
timeout 600 python tools/prof_step.py horse 3 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_horse.csv python tools/prof_step.py horse 3 > gpurun_out/ncu_launch.log 2>&1
# the launch list of the bench command itself (timed region only: skip set-up + 3 warm-up steps)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 470 -c 330 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launch_bench.log 2>&1
if [ -n "$1" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$1" -s 2 -c 1 -o gpurun_out/full_$1 python tools/prof_step.py horse 3 > gpurun_out/ncu_full.log 2>&1
fi
tail -3 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log; cat gpurun_out/quick.log; cat gpurun_out/cppbench_cow.txt
