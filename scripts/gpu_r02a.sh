#!/bin/bash
# round 2, call A: the new parity tests, both bench arms, both CFG forms, kernel microbench baseline.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 1700 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -n 40 gpurun_out/tests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.json | cut -c1-400; tail -n 5 gpurun_out/bench.err
python bench.py --steps 10 --warmup 3 --cfg-mode two-pass --no-cpu-baseline > gpurun_out/bench_twopass.json 2> gpurun_out/bench_twopass.err; echo "bench twopass rc=$?"; tail -n 1 gpurun_out/bench_twopass.json | cut -c1-300
VST_MERGE_LORA=0 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lowrank.json 2> /dev/null; echo "bench lowrank rc=$?"; tail -n 1 gpurun_out/bench_lowrank.json | cut -c1-300
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"; tail -n 1 gpurun_out/bench_reference.json | cut -c1-300
python scripts/kbench.py layout norm tattn gemm gemmres attn bwd > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
ls -la gpurun_out | tail -n 20
