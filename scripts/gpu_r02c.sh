#!/bin/bash
# round 2, call C: residual-epilogue buffer depth A/B, ncu of the L0 residual GEMM, torch-eager context line.
mkdir -p gpurun_out
python scripts/kbench.py gemmres > gpurun_out/kbench_c_4_2.log 2>&1; echo "kbench rc=$?"; grep "+res" gpurun_out/kbench_c_4_2.log | grep -v same
VST_LIB_PATH=$PWD/video_style_transfer_b200/libvst_b200_alt.so python scripts/kbench.py gemmres > gpurun_out/kbench_c_6_3.log 2>&1; echo "kbench alt rc=$?"; grep "+res" gpurun_out/kbench_c_6_3.log | grep -v same
python scripts/torch_eager_context.py --steps 5 --warmup 2 > gpurun_out/torch_eager_context.json 2> gpurun_out/torch_eager_context.err; echo "eager ctx rc=$?"; cat gpurun_out/torch_eager_context.json; tail -n 3 gpurun_out/torch_eager_context.err
KBENCH_ITERS=2 KBENCH_WARMUP=1 python scripts/kbench.py gemmres > gpurun_out/plain_res.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:lora_gemm2_kernel -c 6 -o gpurun_out/prof_res python scripts/kbench.py gemmres > gpurun_out/ncu_res.log 2>&1
echo "ncu res rc=$?"
if [ -f gpurun_out/prof_res.ncu-rep ]; then
  ncu -i gpurun_out/prof_res.ncu-rep --page raw --csv > gpurun_out/prof_res_raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_res.ncu-rep --page source --csv > gpurun_out/prof_res_source.csv 2>/dev/null
  rm -f gpurun_out/prof_res.ncu-rep; gzip -f gpurun_out/prof_res_source.csv
fi
ls -la gpurun_out | tail
