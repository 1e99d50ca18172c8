#!/bin/bash
# Round 2, final 1-GPU verification of the committed tree: the whole GPU suite, smoke, the driver's two bench commands with
# its own arguments (wall time noted), cow, the C++ harnesses, the checked build, the ncu launch list of bench.py.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out /tmp/cl
O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
( time timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/bench_reference.json 2> $O/bench_reference.err ) 2> $O/bench_reference.time
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_default.json 2> $O/bench_default.err ) 2> $O/bench_default.time
timeout 300 python bench.py --workload cow --steps 20 --warmup 5 > $O/wl_cow.json 2> $O/wl_cow.err
for f in cow_ref cow_tr1 horse_ref horse_tr1; do xz -dc tests/golden/clouds/$f.txt.xz > /tmp/cl/$f.txt; done
iterative-closest-point_b200/bin/bench /tmp/cl/cow_ref.txt /tmp/cl/cow_tr1.txt > $O/cppbench_cow.txt 2>/dev/null
iterative-closest-point_b200/bin/bench /tmp/cl/horse_ref.txt /tmp/cl/horse_tr1.txt > $O/cppbench_horse.txt 2>/dev/null
bash tools/gpu_refsrc_check.sh > /dev/null 2>&1
bash tools/checked_run.sh > $O/checked_run.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_bench.csv python bench.py --workload horse --steps 2 --warmup 3 --skip-cpu-baseline > $O/ncu_launch_bench.log 2>&1
tail -n 4 $O/pytest_gpu.log; tail -n 1 $O/smoke.log; cat $O/bench_default.time $O/bench_reference.time | grep real; cut -c1-260 $O/bench_default.json $O/bench_reference.json $O/wl_cow.json
grep -E "opti_gpu_loop|resident_loop" $O/cppbench_cow.txt $O/cppbench_horse.txt; grep -E "loop|IDENT" $O/refsrc_check.txt; cat $O/checked_run.log
