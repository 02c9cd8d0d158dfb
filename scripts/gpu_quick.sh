#!/bin/bash
# Short GPU call: parity suite + headline bench + attention / GEMM microbench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/tests.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.json | cut -c1-400
python scripts/kbench.py attn gemm > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"; tail -n 30 gpurun_out/kbench.log
