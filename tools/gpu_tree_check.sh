#!/bin/bash
# One gpurun call for the tree search: its GPU tests in both forms, bench lines, ncu launch list of a tree-mode loop and
# one full capture per search kernel.   usage: bash tools/gpu_tree_check.sh
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
export ICP_NN=tree
timeout 900 python -m pytest tests/test_gpu_tree.py -m gpu -q > gpurun_out/pytest_tree.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tree.log
ICP_TREE=binary timeout 900 python -m pytest tests/test_gpu_tree.py -m gpu -q > gpurun_out/pytest_tree_binary.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tree_binary.log
for w in cow horse horse_tr2 bunny syn64k syn256k syn1m syn4m; do
  timeout 600 python bench.py --nn tree --skip-cpu-baseline --workload $w --steps 10 --warmup 3 > gpurun_out/tree_wl_$w.json 2> gpurun_out/tree_wl_$w.err
done
timeout 600 python tools/prof_step.py horse 30 > gpurun_out/tree_prof_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_tree_horse.csv python tools/prof_step.py horse 30 > gpurun_out/ncu_tree_launch.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:nn_wtree" -s 12 -c 1 -o gpurun_out/full_nn_wtree python tools/prof_step.py horse 30 > gpurun_out/ncu_tree_full_w.log 2>&1
timeout 600 python tools/prof_step.py syn1m 3 > gpurun_out/tree_prof_plain_1m.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:nn_tree_kernel" -s 2 -c 1 -o gpurun_out/full_nn_tree python tools/prof_step.py syn1m 3 > gpurun_out/ncu_tree_full_b.log 2>&1
tail -2 gpurun_out/pytest_tree.log gpurun_out/pytest_tree_binary.log
for w in cow horse horse_tr2 bunny syn64k syn256k syn1m syn4m; do python - <<PY
import json
d = json.loads(open("gpurun_out/tree_wl_$w.json").read().strip().splitlines()[-1])
print("$w", "ms/step %.3f" % d["ms_per_step"], "it/s %.1f" % d["icp_iterations_per_sec"], "iters", d["iterations_per_step"], "launches", d["gpu_launches"], "e2e ms %.3f" % d["e2e"]["ms_per_step"])
PY
done
