set -x
python scripts/dump_forward.py gpurun_out/ps0.npz > gpurun_out/ps0.log 2>&1
GB_PAIR_PER_SAMPLE=1 python scripts/dump_forward.py gpurun_out/ps1.npz > gpurun_out/ps1.log 2>&1
python - <<'PY' > gpurun_out/ps_cmp.log 2>&1
import numpy as np
a=np.load('gpurun_out/ps0.npz'); b=np.load('gpurun_out/ps1.npz')
for k in a.files: print(k, np.array_equal(a[k],b[k]), float(np.abs(a[k]-b[k]).max()))
PY
python bench.py --no-cpu-baseline > gpurun_out/ps_bench0.log 2>&1
GB_PAIR_PER_SAMPLE=1 python bench.py --no-cpu-baseline > gpurun_out/ps_bench1.log 2>&1
GB_PAIR_PER_SAMPLE=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/ps_launches1.csv python scripts/kbench.py > gpurun_out/ps_ncu.log 2>&1
cat gpurun_out/ps_cmp.log; grep -o '"ms_per_step": [0-9.]*' gpurun_out/ps_bench0.log gpurun_out/ps_bench1.log
timeout 600 python -m pytest tests/test_gpu_parity.py -q -s -k c1_trajectory > gpurun_out/ps_c1.log 2>&1; grep "C1 \|passed\|failed" gpurun_out/ps_c1.log
