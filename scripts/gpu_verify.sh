#!/bin/bash
# One verification round on the GPU box: parity tests, the contract bench line, per-kernel times of both tensor-core
# modes and the ncu launch list of the headline mode.
# usage (from the repo root, under gpurun): bash scripts/gpu_verify.sh <tag>
tag=${1:-v}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
python bench.py > gpurun_out/${tag}_bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/${tag}_bench.log
GB_KBENCH_PREC=bf16x3 python scripts/kbench.py > gpurun_out/${tag}_kbench_x3.log 2>&1
GB_KBENCH_PREC=bf16 python scripts/kbench.py > gpurun_out/${tag}_kbench_bf16.log 2>&1
GB_KBENCH_PREC=bf16x3 timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${tag}_launches_x3.csv python scripts/kbench.py > gpurun_out/${tag}_ncu.log 2>&1
tail -3 gpurun_out/${tag}_pytest.log; grep -o '"ms_per_step": [0-9.]*' gpurun_out/${tag}_bench.log; cat gpurun_out/${tag}_kbench_x3.log gpurun_out/${tag}_kbench_bf16.log
