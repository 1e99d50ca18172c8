#!/bin/bash
# Round 2, 1-GPU call: ICP_NN_AUTO (the library's default: the tree inside icp_run on clouds too small for one wave of the
# brute-force filter).  Its tests first, then the whole suite, smoke, and what a C++ caller of the drop-in now sees on cow.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out /tmp/cl
O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_tree.py tests/test_gpu_driver.py -m gpu -x -q -k "auto_mode or tree_search_driver" > $O/pytest_auto.log 2>&1; echo "rc=$?" >> $O/pytest_auto.log
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
for f in cow_ref cow_tr1 horse_ref horse_tr1; do xz -dc tests/golden/clouds/$f.txt.xz > /tmp/cl/$f.txt; done
for mode in auto brute tree; do
  ICP_NN=$mode iterative-closest-point_b200/bin/bench /tmp/cl/cow_ref.txt /tmp/cl/cow_tr1.txt > $O/cppbench_cow_$mode.txt 2>/dev/null
done
iterative-closest-point_b200/bin/bench /tmp/cl/horse_ref.txt /tmp/cl/horse_tr1.txt > $O/cppbench_horse_auto.txt 2>/dev/null
bash tools/gpu_refsrc_check.sh > /dev/null 2>&1
for mode in auto brute; do ICP_NN=$mode QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/quick_small_$mode.log 2>&1; done
timeout 300 python bench.py --workload cow --steps 20 --warmup 5 --skip-cpu-baseline > $O/wl_cow.json 2> $O/wl_cow.err
tail -n 4 $O/pytest_auto.log; tail -n 4 $O/pytest_gpu.log; tail -n 1 $O/smoke.log
grep -HE "opti_gpu_loop|resident_loop" $O/cppbench_cow_*.txt $O/cppbench_horse_auto.txt; grep -E "loop|IDENT" $O/refsrc_check.txt; grep -H "loop" $O/quick_small_*.log; cut -c1-200 $O/wl_cow.json
