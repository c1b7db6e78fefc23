#!/bin/bash
# `ncu --set full` captures of the hot kernels at the C2 shape (B=4, L=256), run AFTER the same program has exited 0
# without ncu.  usage: ncu_capture.sh <tag> [precision]   -> gpurun_out/<tag>_{pair,ipa}.ncu-rep + <tag>_{pair,ipa}_raw.csv
# kbench launches 52 pair-stack kernels in its per-kernel warm-ups and 45 per forward: skip them and one warm forward,
# then capture the first layer of the next forward (K4 a+g, K4 b, K5, K6, twice, then K7) in its real context; the
# attention kernel is captured by a second pass (8 launches per forward).  gpurun copies back at most 64 MiB.
tag=${1:-cap}
export GB_KBENCH_PREC=${2:-bf16x3}
rx="trimul_out_x3_kernel|trimul_out_tc_kernel|pair_transition_x3_kernel|pair_transition_tc_kernel|trimul_contract_pers_tc_kernel|trimul_proj_x3_kernel|trimul_proj_tc_kernel"
mkdir -p gpurun_out
python scripts/kbench.py > gpurun_out/${tag}_plain.log 2>&1 || { echo "kbench failed without ncu"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name "regex:${rx}" --launch-skip ${NCU_SKIP:-97} --launch-count ${NCU_COUNT:-9} \
  -f -o gpurun_out/${tag}_pair python scripts/kbench.py > gpurun_out/${tag}_ncu_pair.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name "regex:ipa_attention_tc_kernel|ipa_prep_tc_kernel|pair_embed" --launch-skip 10 --launch-count 4 \
  -f -o gpurun_out/${tag}_ipa python scripts/kbench.py > gpurun_out/${tag}_ncu_ipa.log 2>&1
for k in pair ipa; do ncu -i gpurun_out/${tag}_$k.ncu-rep --page raw --csv > gpurun_out/${tag}_${k}_raw.csv 2>/dev/null; done
ls -la gpurun_out/
