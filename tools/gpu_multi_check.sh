#!/bin/bash
# gpurun --gpus 2: GPU tests incl. the 2-GPU parity test, dist_check for all exchange paths, weak-scaling bench lines.
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out; rm -f gpurun_out/dist_check.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n2.log
for c in cow_tr1 horse_tr2; do timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/dist_check.py $c >> gpurun_out/dist_check.log 2>&1; echo "rc=$?" >> gpurun_out/dist_check.log; done
timeout 300 python bench.py --steps 20 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
ICP_ALIGN=fused timeout 300 python bench.py --steps 20 > gpurun_out/bench_n1_fused.json 2> gpurun_out/bench_n1_fused.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_n2_peer.json 2> gpurun_out/bench_n2_peer.err
ICP_ALIGN=fused timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_n2_peer_fused.json 2> gpurun_out/bench_n2_peer_fused.err
ICP_COMM=nccl timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_n2_nccl.json 2> gpurun_out/bench_n2_nccl.err
ICP_PROFILE=1 timeout 600 python tools/quick_bench.py > gpurun_out/quick_profile_split.log 2>&1
ICP_ALIGN=fused ICP_PROFILE=1 timeout 600 python tools/quick_bench.py > gpurun_out/quick_profile_fused.log 2>&1
tail -3 gpurun_out/pytest_gpu_n2.log; grep -v "^\*\|OMP\|^$" gpurun_out/dist_check.log
head -6 gpurun_out/quick_profile_split.log; head -6 gpurun_out/quick_profile_fused.log
for f in bench_n1 bench_n1_fused bench_n2_peer bench_n2_peer_fused bench_n2_nccl; do python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1])
    print('$f', round(d['value']/1e12,3), 'T evals/s', round(d['ms_per_step'],3), 'ms/step  e2e', round(d['e2e']['value']/1e12,3))
except Exception as e: print('$f', e, open('gpurun_out/$f.err').read()[-600:])
PY
done
