#!/bin/bash
# round 2, evidence set: both bench arms, secondary lines, kbench, ncu launch list of one step, ncu --set full of
# the dominant GEMM and of the attention forward kernel.  scripts/collect_profiles.py r02 turns it into profiles/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"
python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench.err
python bench.py --steps 10 --warmup 3 --cfg-mode two-pass --no-cpu-baseline > gpurun_out/bench_twopass.json 2>/dev/null; echo "two-pass rc=$?"
VST_MERGE_LORA=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lowrank.json 2>/dev/null; echo "lowrank rc=$?"
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --with-spatial-ff > gpurun_out/bench_spatial_ff.json 2>/dev/null; echo "spatial-ff rc=$?"
python bench.py --workload cfg2 --steps 10 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "cfg2 rc=$?"
python bench.py --workload cfg4 --steps 3 > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err; echo "cfg4 rc=$?"
python scripts/torch_eager_context.py --steps 5 --warmup 2 > gpurun_out/torch_eager_context.json 2>/dev/null; echo "eager ctx rc=$?"
python scripts/kbench.py layout norm tattn gemm dom fused gemmres attn bwd > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
python bench.py --ncu-step --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --ncu-step --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
KBENCH_ITERS=2 KBENCH_WARMUP=1 python scripts/kbench.py dom > gpurun_out/plain2.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:lora_gemm2_kernel -c 6 -o gpurun_out/prof_gemm python scripts/kbench.py dom > gpurun_out/ncu_gemm.log 2>&1
echo "ncu gemm rc=$?"
KBENCH_ITERS=2 KBENCH_WARMUP=1 python scripts/kbench.py attn > gpurun_out/plain1.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:attn_fwd -c 8 -o gpurun_out/prof_attn python scripts/kbench.py attn > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn rc=$?"
for n in attn gemm; do
  if [ -f gpurun_out/prof_$n.ncu-rep ]; then
    ncu -i gpurun_out/prof_$n.ncu-rep --page raw --csv > gpurun_out/prof_${n}_raw.csv 2>/dev/null
    ncu -i gpurun_out/prof_$n.ncu-rep --page source --csv > gpurun_out/prof_${n}_source.csv 2>/dev/null
    rm -f gpurun_out/prof_$n.ncu-rep
  fi
done
gzip -f gpurun_out/prof_*_source.csv 2>/dev/null
ls -la gpurun_out | tail -n 30
