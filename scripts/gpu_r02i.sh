#!/bin/bash
# round 2, call I: single-launch fused LoRA GEMM + module-level C ABI -- full parity suite, microbench, low-rank bench lines.
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q > gpurun_out/tests_i.log 2>&1; echo "all tests rc=$?"; tail -n 8 gpurun_out/tests_i.log | cut -c1-300
python scripts/kbench.py fused > gpurun_out/kbench_fused.log 2>&1; echo "kbench rc=$?"; cat gpurun_out/kbench_fused.log
VST_MERGE_LORA=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lowrank_fused1.json 2> gpurun_out/bench_i.err; echo "lowrank (fused=1) rc=$?"; tail -n 3 gpurun_out/bench_i.err
VST_MERGE_LORA=0 VST_FUSED_LORA=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lowrank_fused0.json 2>/dev/null; echo "lowrank (fused=0) rc=$?"
VST_MERGE_LORA=0 VST_FUSED_LORA=2 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lowrank_fused2.json 2>/dev/null; echo "lowrank (fused=2) rc=$?"
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_merged.json 2>/dev/null; echo "merged rc=$?"
python - <<'PY'
import json
for f in ("bench_lowrank_fused1","bench_lowrank_fused0","bench_lowrank_fused2","bench_merged"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, "ms", round(d["ms_per_step"],2), "steps/s", round(d["value"],3), "e2e", round(d["e2e"]["value"],3), "MHz", d["clocks"]["sm_mhz"], "launches", d["gpu_launches"])
    except Exception as e: print(f, "ERR", e)
PY
