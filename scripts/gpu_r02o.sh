#!/bin/bash
# GroupNorm statistics without index divisions, TwoLaunchLoraFn / dx on the CTA-pair GEMM: parity suite,
# norm microbench, cfg4 training step, headline
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/tests_o.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/tests_o.log | cut -c1-250
python scripts/kbench.py norm > gpurun_out/kbench_norm_o.log 2>&1; grep -i "groupnorm" gpurun_out/kbench_norm_o.log
python bench.py --workload cfg4 --steps 3 > gpurun_out/bench_cfg4_o.json 2> gpurun_out/bench_cfg4_o.err; echo "cfg4 rc=$?"; tail -n 1 gpurun_out/bench_cfg4_o.json | cut -c1-400
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_o.json 2> gpurun_out/bench_o.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench_o.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_o.json").read().strip().splitlines()[-1])
print("ms", round(d["ms_per_step"],2), "steps/s", round(d["value"],3), "e2e", round(d["e2e"]["value"],3), "frac_step", round(d["frac_step"],3), d["clocks"])
PY
