#!/bin/bash
# One `ncu --set full` capture of the hot kernels at the C2 shape (B=4, L=256), run AFTER the same program has
# exited 0 without ncu.  usage: ncu_capture.sh <tag> [precision] [kernel regex]   -> gpurun_out/<tag>.ncu-rep + <tag>_raw.csv
# kbench launches 52 matching kernels in its per-kernel warm-ups and 53 per forward: skip them and one warm forward, then
# capture one whole forward's worth (every kernel in its real context).
tag=${1:-cap}
export GB_KBENCH_PREC=${2:-bf16x3}
rx=${3:-"trimul_out_x3_kernel|trimul_out_tc_kernel|pair_transition_x3_kernel|pair_transition_tc_kernel|trimul_contract_pers_tc_kernel|trimul_proj_x3_kernel|trimul_proj_tc_kernel|ipa_attention_tc_kernel"}
mkdir -p gpurun_out
python scripts/kbench.py > gpurun_out/${tag}_plain.log 2>&1 || { echo "kbench failed without ncu"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name "regex:${rx}" --launch-skip ${NCU_SKIP:-105} --launch-count ${NCU_COUNT:-53} \
  -f -o gpurun_out/${tag} python scripts/kbench.py > gpurun_out/${tag}_ncu.log 2>&1
ncu -i gpurun_out/${tag}.ncu-rep --page raw --csv > gpurun_out/${tag}_raw.csv 2>/dev/null
ls -la gpurun_out/${tag}.ncu-rep
