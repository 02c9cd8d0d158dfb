#!/bin/bash
# ncu --set full of the kernels changed last: attention forward (self shapes unchanged, 80-key cross instantiation
# new) and the CTA-pair vst_gemm
mkdir -p gpurun_out
python scripts/kbench.py attn > gpurun_out/plain1.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:attn_fwd -c 8 -o gpurun_out/prof_attn python scripts/kbench.py attn > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn rc=$?"
python scripts/kbench.py bwd > gpurun_out/plain3.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none -k regex:gemm_t2 -c 8 -o gpurun_out/prof_gemmt2 python scripts/kbench.py bwd > gpurun_out/ncu_gemmt2.log 2>&1
echo "ncu gemm_t2 rc=$?"
for n in attn gemmt2; do
  if [ -f gpurun_out/prof_$n.ncu-rep ]; then
    ncu -i gpurun_out/prof_$n.ncu-rep --page raw --csv > gpurun_out/prof_${n}_raw.csv 2>/dev/null
    rm -f gpurun_out/prof_$n.ncu-rep
  fi
done
ls -la gpurun_out | tail -8
