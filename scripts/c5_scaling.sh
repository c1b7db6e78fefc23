#!/bin/bash
# C5 (L=1024, batch 1, row-sharded) at several world sizes on one box; one JSON line per run under gpurun_out/.
# usage: scripts/c5_scaling.sh TAG "8 4 2"   (optionally NCCL_* variables in the environment)
tag=$1; shift
for n in $1; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29550 + n)) \
    bench.py --workload c5 --gpus $n --steps 10 --warmup 3 2>gpurun_out/c5_${tag}_n${n}.err | grep "^{" > gpurun_out/c5_${tag}_n${n}.json
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/c5_${tag}_n${n}.json').read())
    print('${tag} n=${n}: %.2f ms/eval, collectives %s' % (d['ms_per_step'], d.get('collectives')))
except Exception as e:
    print('${tag} n=${n}: failed', e)
PY
done
