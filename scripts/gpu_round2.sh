#!/bin/bash
# One GPU call: full parity suite, headline bench + secondary workloads, kernel microbench,
# ncu launch list of one step, ncu --set full of the GEMM / attention-backward kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/tests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.json | cut -c1-600
python bench.py --workload cfg2 --steps 10 > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "cfg2 rc=$?"; tail -n 1 gpurun_out/bench_cfg2.json | cut -c1-900
python bench.py --workload cfg4 --steps 3 > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err; echo "cfg4 rc=$?"; tail -n 1 gpurun_out/bench_cfg4.json | cut -c1-900; tail -n 5 gpurun_out/bench_cfg4.err
python bench.py --workload cfg5 --steps 3 > gpurun_out/bench_cfg5_p1.json 2> gpurun_out/bench_cfg5_p1.err; echo "cfg5 p1 rc=$?"; tail -n 1 gpurun_out/bench_cfg5_p1.json | cut -c1-600
python scripts/kbench.py layout norm tattn gemm attn bwd > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
python bench.py --ncu-step --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --ncu-step --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
python scripts/kbench.py gemm > gpurun_out/plain2.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:lora_gemm -c 16 -o gpurun_out/prof_gemm python scripts/kbench.py gemm > gpurun_out/ncu_gemm.log 2>&1
echo "ncu gemm rc=$?"
python scripts/kbench.py bwd > gpurun_out/plain3.log 2>&1 &&
KBENCH_ITERS=1 KBENCH_WARMUP=1 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -c 8 -o gpurun_out/prof_attn_bwd python scripts/kbench.py bwd > gpurun_out/ncu_attn_bwd.log 2>&1
echo "ncu attn_bwd rc=$?"
