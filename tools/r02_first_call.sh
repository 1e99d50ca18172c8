#!/bin/bash
# First gpurun call of the next round: regression (GPU tests, smoke, default bench) and the experiments staged at the
# end of round 1 without GPU time: ICP_GRAPH (tools/graph_check.py), ICP_NN_MIN_UNITS (small clouds) and the build-time
# variants of filter / refine / exact (tools/filter_ab.sh).  BEFORE the call, in the build container:  make -C iterative-closest-point_b200 variants
#   gpurun --timeout 3000 -- 'bash tools/r02_first_call.sh'
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_horse.json 2> gpurun_out/bench_horse.err
timeout 600 python tools/graph_check.py > gpurun_out/graph_check.log 2>&1; echo "rc=$?" >> gpurun_out/graph_check.log
ICP_PROFILE=1 timeout 600 python tools/quick_bench.py > gpurun_out/quick_profile.log 2>&1
# small clouds: fewer chunk-units per filter CTA than one staged tile (cow: 18 CTAs -> up to 144); parity first, then the cow lines
for u in 4 2 1; do
  ICP_NN_MIN_UNITS=$u timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/minunits_${u}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/minunits_${u}_pytest.log
  ICP_NN_MIN_UNITS=$u QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > gpurun_out/minunits_${u}_quick.log 2>&1
  ICP_NN_MIN_UNITS=$u timeout 300 python bench.py --skip-cpu-baseline --workload cow --steps 20 --warmup 3 > gpurun_out/minunits_${u}_cow.json 2> gpurun_out/minunits_${u}_cow.err
done
timeout 300 python bench.py --skip-cpu-baseline --workload cow --steps 20 --warmup 3 > gpurun_out/minunits_8_cow.json 2> gpurun_out/minunits_8_cow.err
# tree search: global k-d levels also for models below 131 072 points (emulation: horse 4.70 -> 3.79 leaves per search)
ICP_KD_GLOBAL_MIN=16384 timeout 900 python -m pytest tests/test_gpu_tree.py -m gpu -x -q > gpurun_out/kdmin_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/kdmin_pytest.log
for w in horse horse_tr2 bunny syn64k; do
  timeout 300 python bench.py --nn tree --skip-cpu-baseline --workload $w --steps 10 --warmup 3 > gpurun_out/kdmin_default_$w.json 2>/dev/null
  ICP_KD_GLOBAL_MIN=16384 timeout 300 python bench.py --nn tree --skip-cpu-baseline --workload $w --steps 10 --warmup 3 > gpurun_out/kdmin_16k_$w.json 2>/dev/null
done
timeout 2400 bash tools/filter_ab.sh > gpurun_out/filter_ab.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log; cat gpurun_out/graph_check.log; head -3 gpurun_out/minunits_*_quick.log; tail -2 gpurun_out/minunits_*_pytest.log gpurun_out/kdmin_pytest.log; grep -ho '"ms_per_step": [0-9.]*' gpurun_out/kdmin_*.json | paste - - ; cat gpurun_out/ab_summary.txt
