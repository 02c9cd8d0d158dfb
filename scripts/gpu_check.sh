#!/bin/bash
# Runs the GPU parity tests group by group (one process each, so a trapped kernel cannot poison
# the next group) and collects logs under gpurun_out/.  Usage: gpu_check.sh [group ...]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
groups=("rearrange" "groupnorm or layernorm" "temporal_attn" "lora_gemm or lora_pack" "attn_fwd"
        "temporal_transformer" "unziplora or lora_compatible or temporal_lora_linear"
        "attn_processor" "motion_module")
if [ $# -gt 0 ]; then groups=("$@"); fi
status=0
: > gpurun_out/summary.txt
for k in "${groups[@]}"; do
  name=$(echo "$k" | tr ' ' '_')
  timeout 600 python -m pytest tests -m gpu -q -x -k "$k" > "gpurun_out/test_${name}.log" 2>&1
  rc=$?
  echo "group [$k] rc=$rc" | tee -a gpurun_out/summary.txt
  tail -n 12 "gpurun_out/test_${name}.log"
  [ $rc -ne 0 ] && status=1
done
exit $status
