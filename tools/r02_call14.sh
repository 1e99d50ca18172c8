#!/bin/bash
# Round 2, 8-GPU call (short): the final tree at the driver's largest N - sharded parity in every mode, the group tests at
# 8 devices, and the driver's own command (bench.py --gpus 8, with its parity gate and secondary.horse).
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29542 bench.py --gpus 8 --steps 5 --warmup 3 > $O/bench_default_n8.json 2> $O/bench_default_n8.err; echo "rc=$?" >> $O/bench_default_n8.err
timeout 120 $TR --master-port 29541 tools/dist_check.py horse_tr2 cow_tr1 > $O/dist_check_n8.raw 2>&1; echo "rc=$?" >> $O/dist_check_n8.raw
grep -E "^dist_check|^rc=|Error|error" $O/dist_check_n8.raw > $O/dist_check_n8.log
timeout 120 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "group_sharded_parity and 8-" > $O/pytest_group_n8.log 2>&1; echo "rc=$?" >> $O/pytest_group_n8.log
grep -c "parity ok" $O/dist_check_n8.log; grep -v "parity ok" $O/dist_check_n8.log; tail -n 3 $O/pytest_group_n8.log; cut -c1-300 $O/bench_default_n8.json; tail -n 1 $O/bench_default_n8.err
