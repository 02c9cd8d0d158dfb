#!/bin/bash
# One GPU call refreshing every number kept under profiles/: parity suite, headline bench (both
# arms), secondary workloads, kernel microbench, ncu launch list, ncu --set full of the attention
# forward / GEMM / attention backward kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/tests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.json | cut -c1-300
VST_MERGE_LORA=0 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lowrank.json 2> /dev/null; echo "bench lowrank rc=$?"; tail -n 1 gpurun_out/bench_lowrank.json | cut -c1-300
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"; tail -n 1 gpurun_out/bench_reference.json | cut -c1-300
python bench.py --workload cfg2 --steps 10 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "cfg2 rc=$?"
python bench.py --workload cfg4 --steps 3 > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err; echo "cfg4 rc=$?"
python bench.py --workload cfg5 --steps 3 > gpurun_out/bench_cfg5_p1.json 2> gpurun_out/bench_cfg5_p1.err; echo "cfg5 p1 rc=$?"
python scripts/kbench.py layout norm tattn gemm gemmres attn bwd > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
python bench.py --ncu-step --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --ncu-step --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
python scripts/kbench.py attn > gpurun_out/plain1.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:attn_fwd -c 8 -o gpurun_out/prof_attn python scripts/kbench.py attn > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn rc=$?"
python scripts/kbench.py gemm > gpurun_out/plain2.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:lora_gemm -c 16 -o gpurun_out/prof_gemm python scripts/kbench.py gemm > gpurun_out/ncu_gemm.log 2>&1
echo "ncu gemm rc=$?"
python scripts/kbench.py bwd > gpurun_out/plain3.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -c 8 -o gpurun_out/prof_attn_bwd python scripts/kbench.py bwd > gpurun_out/ncu_attn_bwd.log 2>&1
echo "ncu attn_bwd rc=$?"
# keep gpurun_out small (64 MiB cap): export the raw pages as CSV on the box, drop the .ncu-rep files
for n in attn gemm attn_bwd; do
  if [ -f gpurun_out/prof_$n.ncu-rep ]; then
    ncu -i gpurun_out/prof_$n.ncu-rep --page raw --csv > gpurun_out/prof_${n}_raw.csv 2>/dev/null
    ncu -i gpurun_out/prof_$n.ncu-rep --page source --csv > gpurun_out/prof_${n}_source.csv 2>/dev/null
    rm -f gpurun_out/prof_$n.ncu-rep
  fi
done
gzip -f gpurun_out/prof_*_source.csv 2>/dev/null
ls -la gpurun_out | tail -n 40
