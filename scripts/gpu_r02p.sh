#!/bin/bash
# lane-owned LayerNorm: parity, A/B against the flat kernels
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "layernorm" > gpurun_out/tests_p.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/tests_p.log | cut -c1-250
for m in ${LN_MODES:-0 1 3 5}; do echo "== VST_LN_LANE=$m"; VST_LN_LANE=$m python scripts/ln_probe.py 2>&1 | grep -v checksum; done > gpurun_out/ln_probe_p.log 2>&1
grep -E "==|gamma\+beta\+pe|gamma\+beta\]" gpurun_out/ln_probe_p.log | grep -E "==|L1"
