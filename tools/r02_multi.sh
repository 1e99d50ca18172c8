#!/bin/bash
# Round 2 multi-GPU evidence on N GPUs of one box (gpurun --gpus N -- 'bash tools/r02_multi.sh N'): sharded parity in
# every exchange mode and both search modes (dist_check), the icp_group_* tests at this device count, bench.py --gpus N
# (the driver's command: 2^22 synthetic cloud, strong scaling, parity gate), config 5 at 1 024 pairs, the tree search,
# and the horse-sized weak probe with both alignment forms.
N=${1:-8}
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/topo_n$N.txt 2>&1
timeout 900 $TR --master-port 29521 tools/dist_check.py horse_tr2 cow_tr1 > $O/dist_check_n$N.raw 2>&1; echo "rc=$?" >> $O/dist_check_n$N.raw
grep -E "^dist_check|^rc=|Error|error" $O/dist_check_n$N.raw > $O/dist_check_n$N.log
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "group_sharded_parity and ${N}-" > $O/pytest_group_n$N.log 2>&1; echo "rc=$?" >> $O/pytest_group_n$N.log
timeout 900 $TR --master-port 29522 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_default_n$N.json 2> $O/bench_default_n$N.err; echo "rc=$?" >> $O/bench_default_n$N.err
timeout 900 $TR --master-port 29523 bench.py --gpus $N --workload batch64k --pairs 1024 --steps 2 --warmup 1 > $O/batch64k_1024_n$N.json 2> $O/batch64k_1024_n$N.err
timeout 600 $TR --master-port 29524 bench.py --gpus $N --nn tree --steps 10 --warmup 3 --no-secondary --skip-parity > $O/tree_syn4m_n$N.json 2> $O/tree_syn4m_n$N.err
timeout 600 $TR --master-port 29525 bench.py --gpus $N --workload horse --steps 20 --warmup 5 --scaling weak --skip-parity > $O/horse_weak_split_n$N.json 2> $O/horse_weak_split_n$N.err
ICP_ALIGN=fused timeout 600 $TR --master-port 29526 bench.py --gpus $N --workload horse --steps 20 --warmup 5 --scaling weak --skip-parity > $O/horse_weak_fused_n$N.json 2> $O/horse_weak_fused_n$N.err
cat $O/dist_check_n$N.log; tail -n 3 $O/pytest_group_n$N.log
for f in bench_default batch64k_1024 tree_syn4m horse_weak_split horse_weak_fused; do echo "== $f"; cut -c1-420 $O/${f}_n$N.json; tail -n 2 $O/${f}_n$N.err; done
