#!/bin/bash
# round 2, call B: TMA residual epilogue -- parity, gemmres microbench, headline bench.
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -n 15 gpurun_out/tests.log
python scripts/kbench.py gemmres gemm > gpurun_out/kbench_b.log 2>&1; echo "kbench rc=$?"; grep -A1 "res" gpurun_out/kbench_b.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_b.json 2> gpurun_out/bench_b.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench_b.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_b.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["frac_step"], d["e2e"]["value"], d["clocks"])
for k,v in d["by_kernel"].items(): print("  ",k,v)
PY
