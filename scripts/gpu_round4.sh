#!/bin/bash
# End-of-round-2 evidence refresh on the final kernel set: parity suite, headline bench (both arms), secondary
# workloads, kernel microbench, ncu launch list of one step, ncu --set full of the memory-bound kernels changed last
# (lane-owned LayerNorm, GroupNorm statistics).  The attention / GEMM forward kernels are unchanged since the captures
# under profiles/ncu_{attn,gemm}_full_r02.csv.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/tests.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"; tail -n 1 gpurun_out/bench_reference.json | cut -c1-300
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.json | cut -c1-300
VST_MERGE_LORA=0 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lowrank.json 2> /dev/null; echo "bench lowrank rc=$?"
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --with-spatial-ff > gpurun_out/bench_spatial_ff.json 2> /dev/null; echo "bench spatial-ff rc=$?"
python bench.py --workload cfg2 --steps 10 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "cfg2 rc=$?"
VST_KERNEL_TABLE=gpurun_out/cfg4_kernels.tsv python bench.py --workload cfg4 --steps 3 > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err; echo "cfg4 rc=$?"
python scripts/kbench.py layout norm tattn gemm gemmres attn bwd > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
python bench.py --ncu-step --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --ncu-step --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
python scripts/ln_probe.py > gpurun_out/ln_probe.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:layernorm_lane -c 4 -o gpurun_out/prof_ln python scripts/ln_probe.py > gpurun_out/ncu_ln.log 2>&1
echo "ncu ln rc=$?"
python scripts/kbench.py norm > gpurun_out/plain_norm.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:gn_stats -c 4 -o gpurun_out/prof_gn python scripts/kbench.py norm > gpurun_out/ncu_gn.log 2>&1
echo "ncu gn rc=$?"
for n in ln gn; do
  if [ -f gpurun_out/prof_$n.ncu-rep ]; then
    ncu -i gpurun_out/prof_$n.ncu-rep --page raw --csv > gpurun_out/prof_${n}_raw.csv 2>/dev/null
    rm -f gpurun_out/prof_$n.ncu-rep
  fi
done
ls -la gpurun_out | tail -n 40
