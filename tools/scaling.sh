#!/bin/bash
# Scaling lines for DESIGN.md / BASELINE.md: run under `gpurun --gpus G -- bash tools/scaling.sh N1 N2 ...`.
#   horse    weak scaling  (every rank holds one horse-sized scene shard; bench.py default workload)
#   syn4m    strong scaling (one 2^22-point scene cut into N blocks, model replicated), max_iter 2
#   batch64k pair-per-GPU  (independent 65 536-point pairs, 16 per GPU)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
for N in "$@"; do
  if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N"; fi
  timeout 600 $L bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/scale_horse_n$N.json 2> gpurun_out/scale_horse_n$N.err
  timeout 900 $L bench.py --gpus $N --steps 2 --warmup 3 --workload syn4m --scaling strong --max-iter 2 > gpurun_out/scale_syn4m_n$N.json 2> gpurun_out/scale_syn4m_n$N.err
  timeout 900 $L bench.py --gpus $N --steps 2 --warmup 1 --workload batch64k --pairs $((16 * N)) > gpurun_out/scale_batch64k_n$N.json 2> gpurun_out/scale_batch64k_n$N.err
done
for f in gpurun_out/scale_*.json; do python - <<PY
import json
try:
    d=json.loads(open('$f').read().strip().splitlines()[-1])
    print('$f'.split('/')[-1], 'N', d['n_gpus'], round(d['value']/1e12,3), 'T evals/s', round(d['ms_per_step'],3), 'ms/step  e2e', round(d['e2e']['value']/1e12,3), (d.get('roofline') or {}).get('frac'))
except Exception as e: print('$f', 'FAILED', e, open('$f'.replace('.json','.err')).read()[-500:])
PY
done
