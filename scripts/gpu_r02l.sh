#!/bin/bash
# round 2, call L: GEGLU epilogue on packed f32x2 math -- parity, per-level timing, headline bench.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_modules_gpu.py tests/test_parity_full_gpu.py -m gpu -q -x > gpurun_out/tests_l.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/tests_l.log | cut -c1-250
python scripts/kbench.py geglu > gpurun_out/kbench_geglu.log 2>&1; echo "kbench rc=$?"; cat gpurun_out/kbench_geglu.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_l.json 2> gpurun_out/bench_l.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench_l.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_l.json").read().strip().splitlines()[-1])
print("ms", round(d["ms_per_step"],2), "steps/s", round(d["value"],3), "e2e", round(d["e2e"]["value"],3), "frac_step", round(d["frac_step"],3), d["clocks"])
for k,v in d["by_kernel"].items(): print("   ",k,v)
PY
