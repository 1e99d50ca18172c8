#!/bin/bash
# Round 2, 2-GPU call: the multi-GPU paths again after the last two changes (matched points recorded by the search kernels,
# ICP_NN_AUTO): the whole GPU suite on two GPUs, dist_check in every mode, bench.py --gpus 2 with its parity gate.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu_n2.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_n2.log
timeout 900 $TR --master-port 29531 tools/dist_check.py horse_tr2 cow_tr1 > $O/dist_check_n2.raw 2>&1; echo "rc=$?" >> $O/dist_check_n2.raw
grep -E "^dist_check|^rc=|Error|error" $O/dist_check_n2.raw > $O/dist_check_n2.log
timeout 900 $TR --master-port 29532 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_default_n2.json 2> $O/bench_default_n2.err; echo "rc=$?" >> $O/bench_default_n2.err
timeout 600 $TR --master-port 29533 bench.py --gpus 2 --nn tree --steps 10 --warmup 3 --no-secondary --skip-parity > $O/tree_syn4m_n2.json 2> $O/tree_syn4m_n2.err
tail -n 4 $O/pytest_gpu_n2.log; cat $O/dist_check_n2.log; cut -c1-300 $O/bench_default_n2.json $O/tree_syn4m_n2.json; tail -n 1 $O/bench_default_n2.err
