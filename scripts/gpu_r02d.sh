#!/bin/bash
# round 2, call D: streaming-softmax attention kernel -- parity, A/B against the exact kernel, poly split.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_attn_stream_gpu.py tests/test_kernels_gpu.py tests/test_backward_gpu.py -m gpu -q -x > gpurun_out/tests_d.log 2>&1; echo "tests rc=$?"; tail -n 12 gpurun_out/tests_d.log
for p in 2 3 4 1; do echo "== stream poly $p"; VST_ATTN_POLY=$p timeout 300 python scripts/kbench.py attn 2>&1 | grep "attn_fwd"; done
echo "== exact kernel"; VST_ATTN_STREAM=0 timeout 300 python scripts/kbench.py attn 2>&1 | grep -A1 "attn_fwd"
