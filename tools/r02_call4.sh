#!/bin/bash
# Round 2, fourth gpurun call (--gpus 2): the multi-GPU paths on hardware - GPU tests (sharded parity in every exchange
# mode, icp_group_*, the two-GPU driver), bench.py --gpus 2 with its parity gate, dist_check logs, bin/bench --devices,
# config 5 at 1 024 pairs on two GPUs; plus the cow diagnostics and the L2-fetch experiment (one GPU, seconds each).
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out /tmp/cl
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/topo2.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_n2.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_n2.log
timeout 900 $TR --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_default_n2.json 2> $O/bench_default_n2.err; echo "rc=$?" >> $O/bench_default_n2.err
for nn in brute tree; do for c in cow_tr1 horse_tr2; do
  ICP_NN=$nn timeout 600 $TR --master-port 29512 tools/dist_check.py $c >> $O/dist_check_n2.log 2>&1; echo "[$nn $c] rc=$?" >> $O/dist_check_n2.log
done; done
for f in cow_ref cow_tr1 horse_ref horse_tr1; do xz -dc tests/golden/clouds/$f.txt.xz > /tmp/cl/$f.txt; done
iterative-closest-point_b200/bin/bench /tmp/cl/horse_ref.txt /tmp/cl/horse_tr1.txt --devices 0,1 > $O/cppbench_horse_devices01.txt 2>&1
iterative-closest-point_b200/bin/bench /tmp/cl/cow_ref.txt /tmp/cl/cow_tr1.txt --devices 0,1 > $O/cppbench_cow_devices01.txt 2>&1
timeout 900 $TR --master-port 29513 bench.py --gpus 2 --workload batch64k --pairs 1024 --steps 2 --warmup 1 > $O/batch64k_1024_n2.json 2> $O/batch64k_1024_n2.err
timeout 600 $TR --master-port 29514 bench.py --gpus 2 --workload horse --steps 20 --warmup 5 --scaling weak > $O/bench_horse_weak_n2.json 2> $O/bench_horse_weak_n2.err
# ---- one-GPU side experiments ----------------------------------------------------------------------------
V=$PWD/iterative-closest-point_b200/lib/variants
for v in default solve1 ldcv; do
  lib=$V/libicp_b200_$v.so; [ $v = default ] && lib=$PWD/iterative-closest-point_b200/lib/libicp_b200.so
  ICP_B200_LIB=$lib ICP_PROFILE=1 QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/cowdiag_$v.log 2>&1
  ICP_B200_LIB=$lib QB_ONLY_SMALL=1 timeout 300 python tools/quick_bench.py > $O/cowdiag_${v}_noprofile.log 2>&1
done
for g in 32 64 128; do ICP_L2_FETCH=$g ICP_NN=tree timeout 300 python tools/prof_tail.py syn4m 3 > $O/l2fetch_$g.log 2>&1; done
timeout 1200 bash tools/filter_ab.sh r12c3g128u8 r16c2g256u8 fused2 > $O/filter_ab.log 2>&1
bash tools/checked_run.sh > $O/checked_run.log 2>&1
tail -n 4 $O/pytest_gpu_n2.log; cut -c1-1200 $O/bench_default_n2.json; cat $O/dist_check_n2.log | grep -v "^W\|^\*\*\*\|OMP_NUM" ; cat $O/cppbench_horse_devices01.txt | tail -6
cut -c1-500 $O/batch64k_1024_n2.json $O/bench_horse_weak_n2.json; head -4 $O/cowdiag_*.log; cat $O/l2fetch_*.log; cat $O/ab_summary.txt; cat $O/checked_run.log
