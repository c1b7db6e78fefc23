#!/bin/bash
# Quick A/B round on the GPU box: tensor-core parity tests, per-kernel times of both modes, launch list of the headline mode.
# usage: bash scripts/gpu_quick.sh <tag> [pytest files...]
tag=${1:-q}; shift
mkdir -p gpurun_out
tests=${@:-"tests/test_gpu_x3.py tests/test_gpu_tensorcore.py tests/test_gpu_headline.py"}
timeout 900 python -m pytest $tests -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
GB_KBENCH_PREC=bf16x3 python scripts/kbench.py > gpurun_out/${tag}_kbench_x3.log 2>&1
GB_KBENCH_PREC=bf16 python scripts/kbench.py > gpurun_out/${tag}_kbench_bf16.log 2>&1
GB_KBENCH_PREC=bf16x3 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/${tag}_launches_x3.csv python scripts/kbench.py > gpurun_out/${tag}_ncu.log 2>&1
python scripts/summarize_launches.py gpurun_out/${tag}_launches_x3.csv last > gpurun_out/${tag}_launches_x3.txt 2>&1
tail -3 gpurun_out/${tag}_pytest.log; cat gpurun_out/${tag}_kbench_x3.log gpurun_out/${tag}_kbench_bf16.log; head -16 gpurun_out/${tag}_launches_x3.txt
