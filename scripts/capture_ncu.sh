#!/bin/bash
# One gpurun call: every profile target first plain (must exit 0), then under `ncu --set full`; reports land in
# gpurun_out/ncu/<target>.ncu-rep. Afterwards, on the CPU box: python scripts/make_profiles.py
#   gpurun --timeout 1500 -- 'bash scripts/capture_ncu.sh gemm_fold_qkv gemm_fold_ffn1 gemm_fold_out gemm_fold_ffn2 attention'
# Kernel-name filters keep ncu to the launches that matter (it replays each profiled launch ~40 times).
set -u
mkdir -p gpurun_out/ncu
declare -A FILTER=(
  [gemm_ffn1]=gemm_bf16_tn_kernel [gemm_ffn2]=gemm_bf16_tn_kernel [gemm_qkv]=gemm_bf16_tn_kernel [gemm_out]=gemm_bf16_tn_kernel
  [gemm_fold_ffn1]=gemm_bf16_tn_kernel [gemm_fold_ffn2]=gemm_bf16_tn_kernel [gemm_fold_qkv]=gemm_bf16_tn_kernel
  [gemm_fold_out]=gemm_bf16_tn_kernel [attention]=bert_attention_tc [gemm_grad]=gemm_grad_kernel [layernorm]=layernorm_kernel
  [aggregate]=agg_ [ue_attention]=ue_attention_tc_kernel [ue_attention_bwd]=ue_attention_bwd_tc_kernel)
declare -A COUNT=([aggregate]=7)
declare -A SKIP=([aggregate]=14)
for t in "$@"; do
  f=${FILTER[$t]:-}
  if [ -z "$f" ]; then echo "unknown target $t"; continue; fi
  if timeout 120 python scripts/profile_target.py "$t" > "gpurun_out/ncu/plain_$t.log" 2>&1; then
    timeout 400 ncu --set full --clock-control none --import-source on -k "regex:$f" -s ${SKIP[$t]:-2} -c ${COUNT[$t]:-1} -f \
      -o "gpurun_out/ncu/$t" python scripts/profile_target.py "$t" > "gpurun_out/ncu/ncu_$t.log" 2>&1
    echo "$t: ncu exit $?"
  else
    echo "$t: plain run failed, not profiled"; tail -3 "gpurun_out/ncu/plain_$t.log"
  fi
done
