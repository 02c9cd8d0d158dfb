#!/bin/bash
# round 2, call G (2 or more GPUs): sequence-parallel path on real ranks + bench.py under torchrun with the sp probe.
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus.txt 2>&1
timeout 900 python -m pytest tests/test_sp_gpu.py -m gpu -q > gpurun_out/tests_sp.log 2>&1; echo "sp tests rc=$?"; tail -n 6 gpurun_out/tests_sp.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_dp$N.json 2> gpurun_out/bench_dp$N.err; echo "bench rc=$?"; tail -n 5 gpurun_out/bench_dp$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_dp$N.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"])
print(json.dumps(d.get("sp"), indent=1))
PY
