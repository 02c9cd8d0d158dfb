#!/bin/bash
# lane-owned LayerNorm as the default: full parity suite, norm microbench, headline
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/tests_q.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/tests_q.log | cut -c1-250
python scripts/kbench.py norm > gpurun_out/kbench_norm_q.log 2>&1; cat gpurun_out/kbench_norm_q.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench_q.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_q.json").read().strip().splitlines()[-1])
print("ms", round(d["ms_per_step"],2), "steps/s", round(d["value"],3), "e2e", round(d["e2e"]["value"],3), "frac_step", round(d["frac_step"],3), d["clocks"])
PY
