#!/bin/bash
# Runs the GPU parity tests group by group (one process each, so a trapped kernel cannot poison
# the next group) and collects logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
python -c "import diffusers; print('diffusers', diffusers.__version__)" > gpurun_out/diffusers.txt 2>&1
status=0
for k in "rearrange" "groupnorm or layernorm" "temporal_attn" "lora_gemm or lora_pack" "attn_fwd" ${EXTRA_GROUPS}; do
  name=$(echo "$k" | tr ' ' '_')
  timeout 600 python -m pytest tests -m gpu -q -x -k "$k" > "gpurun_out/test_${name}.log" 2>&1
  rc=$?
  echo "group [$k] rc=$rc" | tee -a gpurun_out/summary.txt
  tail -n 15 "gpurun_out/test_${name}.log"
  [ $rc -ne 0 ] && status=1
done
exit $status
