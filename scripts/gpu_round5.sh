#!/bin/bash
# Last refresh of round 2 after the cross-attention work: parity suite, headline (both arms), kernel microbench,
# ncu launch list of one step.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/tests.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference rc=$?"
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 1 gpurun_out/bench.json | cut -c1-300
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --cfg-mode two-pass > gpurun_out/bench_twopass.json 2> /dev/null; echo "bench two-pass rc=$?"
python scripts/kbench.py layout norm tattn gemm gemmres attn bwd > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
python bench.py --ncu-step --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --ncu-step --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
