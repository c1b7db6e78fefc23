#!/bin/bash
# One `ncu --set full` capture of the hot kernels at the C2 shape (B=4, L=256, bf16), run AFTER the same program has
# exited 0 without ncu.  usage: ncu_capture.sh <tag> [kernel regex]      -> gpurun_out/<tag>.ncu-rep + <tag>_raw.csv
tag=${1:-cap}
rx=${2:-"trimul_out_tc_kernel|pair_transition_tc_kernel|trimul_contract_tc_kernel|trimul_proj_tc_kernel|ipa_attention_tc_kernel"}
mkdir -p gpurun_out
python scripts/kbench.py > gpurun_out/${tag}_plain.log 2>&1 || { echo "kbench failed without ncu"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
# the whole forward at the end of kbench holds every kernel in its real context: skip the per-kernel warm-ups
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name "regex:${rx}" --launch-skip ${NCU_SKIP:-68} --launch-count ${NCU_COUNT:-12} \
  -f -o gpurun_out/${tag} python scripts/kbench.py > gpurun_out/${tag}_ncu.log 2>&1
ncu -i gpurun_out/${tag}.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ls -la gpurun_out/${tag}.ncu-rep
