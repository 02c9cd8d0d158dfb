#!/bin/bash
# text cross-attention with an 80-key block: parity, microbench A/B, headline
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_modules_gpu.py tests/test_backward_gpu.py tests/test_parity_full_gpu.py -m gpu -q -x > gpurun_out/tests_n.log 2>&1; echo "tests rc=$?"; tail -n 4 gpurun_out/tests_n.log | cut -c1-250
echo "== 80-key block"; python scripts/kbench.py attn 2>&1 | grep "attn_fwd"
echo "== 128-key block"; VST_ATTN_BKV80=0 python scripts/kbench.py attn 2>&1 | grep "attn_fwd.*cross"
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n.json 2> gpurun_out/bench_n.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench_n.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n.json").read().strip().splitlines()[-1])
print("ms", round(d["ms_per_step"],2), "steps/s", round(d["value"],3), "e2e", round(d["e2e"]["value"],3), "frac_step", round(d["frac_step"],3), d["clocks"])
for k,v in d["by_kernel"].items(): print("   ",k,v)
PY
