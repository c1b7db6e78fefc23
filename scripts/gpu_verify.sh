#!/bin/bash
# One verification round on the GPU box: parity tests, the contract bench line, the length sweep and the launch list.
# usage (from the repo root, under gpurun): bash scripts/gpu_verify.sh <tag>
tag=${1:-v}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
python bench.py > gpurun_out/${tag}_bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/${tag}_bench.log
python scripts/length_sweep.py > gpurun_out/${tag}_sweep5.json 2> gpurun_out/${tag}_sweep.err
python scripts/length_sweep.py 8 > gpurun_out/${tag}_sweep8.json 2>> gpurun_out/${tag}_sweep.err
python scripts/kbench.py > gpurun_out/${tag}_kbench.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${tag}_launches.csv python scripts/kbench.py > gpurun_out/${tag}_ncu.log 2>&1
tail -3 gpurun_out/${tag}_pytest.log; grep -o '"ms_per_step": [0-9.]*' gpurun_out/${tag}_bench.log; cat gpurun_out/${tag}_kbench.log gpurun_out/${tag}_sweep5.json
