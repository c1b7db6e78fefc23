#!/bin/bash
# A/B of one environment switch on the GPU box: bit-compare the forward, time the kernels, run the tensor-core tests.
# usage: gpu_ab.sh VAR   (runs VAR=0 and VAR=1)
var=$1
mkdir -p gpurun_out
env $var=0 python scripts/dump_forward.py gpurun_out/ab0.npz > gpurun_out/ab0.log 2>&1
env $var=1 python scripts/dump_forward.py gpurun_out/ab1.npz > gpurun_out/ab1.log 2>&1
python - <<'PY' > gpurun_out/ab_cmp.log 2>&1
import numpy as np
a=np.load('gpurun_out/ab0.npz'); b=np.load('gpurun_out/ab1.npz')
for k in a.files: print(k, 'bit-identical' if np.array_equal(a[k],b[k]) else 'DIFFERENT', float(np.abs(a[k]-b[k]).max()))
PY
for v in 0 1 0 1; do echo "$var=$v"; env $var=$v python scripts/kbench.py 2>&1 | grep "K5\|K6\|forward"; done > gpurun_out/ab_time.log 2>&1
env $var=1 timeout 600 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_shard.py -x -q > gpurun_out/ab_pytest.log 2>&1
env $var=1 python scripts/kbench.py 5 75 2>&1 | grep "K5\|forward" >> gpurun_out/ab_time.log
cat gpurun_out/ab_cmp.log gpurun_out/ab_time.log; tail -3 gpurun_out/ab_pytest.log; tail -3 gpurun_out/ab1.log
