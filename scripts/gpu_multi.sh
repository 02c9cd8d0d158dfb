#!/bin/bash
# Multi-GPU measurements on N GPUs of one box: SP parity test, cfg5 (one 32-frame video,
# sequence-parallel) and the headline data-parallel line.  Usage: gpu_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "2" ]; then timeout 600 python -m pytest tests/test_sp_gpu.py -m gpu -q -x 2>&1 | tail -3; fi
timeout 900 $TR --master-port 29511 bench.py --workload cfg5 --gpus $N --steps 3 > gpurun_out/bench_cfg5_p$N.json 2> gpurun_out/bench_cfg5_p$N.err
echo "cfg5 rc=$?"; tail -n 1 gpurun_out/bench_cfg5_p$N.json | cut -c1-700; tail -n 3 gpurun_out/bench_cfg5_p$N.err
timeout 900 $TR --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dp$N.json 2> gpurun_out/bench_dp$N.err
echo "dp rc=$?"; tail -n 1 gpurun_out/bench_dp$N.json | cut -c1-400; tail -n 3 gpurun_out/bench_dp$N.err
