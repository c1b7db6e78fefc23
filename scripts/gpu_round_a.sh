mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ga_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ga_pytest.log
GB_STEP_GRAPH=0 python scripts/length_sweep.py > gpurun_out/ga_sweep_plain.json 2> gpurun_out/ga_sweep_plain.err
python scripts/length_sweep.py > gpurun_out/ga_sweep_graph.json 2> gpurun_out/ga_sweep_graph.err
GB_STEP_GRAPH=0 python scripts/length_sweep.py 8 > gpurun_out/ga_sweep8_plain.json 2>> gpurun_out/ga_sweep_plain.err
python scripts/length_sweep.py 8 > gpurun_out/ga_sweep8_graph.json 2>> gpurun_out/ga_sweep_graph.err
python bench.py > gpurun_out/ga_bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/ga_bench.log
NCU_COUNT=14 bash scripts/ncu_capture.sh ga_full > gpurun_out/ga_cap.log 2>&1
tail -3 gpurun_out/ga_pytest.log; cat gpurun_out/ga_sweep*.json; tail -2 gpurun_out/ga_bench.log | cut -c1-600; tail -3 gpurun_out/ga_cap.log
