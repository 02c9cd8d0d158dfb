#!/bin/bash
# One GPU call: parity tests, bench, kernel microbench, ncu launch list, ncu --set full of the top kernel.
mkdir -p gpurun_out
bash scripts/gpu_check.sh > gpurun_out/check.log 2>&1; echo "check rc=$?"
python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench.json
python scripts/kbench.py > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
python bench.py --ncu-step --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --ncu-step --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
python scripts/kbench.py attn > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_fwd -s 3 -c 2 -o gpurun_out/prof_attn python scripts/kbench.py attn > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn rc=$?"
