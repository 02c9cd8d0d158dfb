#!/bin/bash
# Which part of the fused GEMM bounds the small-K shapes: runs the gemm microbench with the
# epilogue stores / the epilogue / the MMAs / the TMA loads removed (VST_GEMM_DEBUG, profiling only).
mkdir -p gpurun_out
for d in ${DIAG:-0 1 2 4 8 12}; do
  echo "== VST_GEMM_DEBUG=$d" >> gpurun_out/gemm_diag.log
  VST_GEMM_DEBUG=$d python scripts/kbench.py gemm 2>&1 | grep "^lora_gemm" >> gpurun_out/gemm_diag.log
done
cat gpurun_out/gemm_diag.log
