#!/bin/bash
# round 2, call E: stream attention -- elected barrier arrivals (main build) vs + ping-pong (alt build).
mkdir -p gpurun_out
ALT=$PWD/video_style_transfer_b200/libvst_b200_alt.so
timeout 600 python -m pytest tests/test_attn_stream_gpu.py tests/test_sp_gpu.py -m gpu -q -x > gpurun_out/tests_e.log 2>&1; echo "tests main rc=$?"; tail -n 4 gpurun_out/tests_e.log
VST_LIB_PATH=$ALT timeout 600 python -m pytest tests/test_attn_stream_gpu.py -m gpu -q -x > gpurun_out/tests_e_alt.log 2>&1; echo "tests alt rc=$?"; tail -n 4 gpurun_out/tests_e_alt.log
for p in 2 3; do echo "== main (elected arrives) poly $p"; VST_ATTN_POLY=$p timeout 300 python scripts/kbench.py attn 2>&1 | grep "attn_fwd"; done
for p in 1 2 3; do echo "== alt (ping-pong) poly $p"; VST_LIB_PATH=$ALT VST_ATTN_POLY=$p timeout 300 python scripts/kbench.py attn 2>&1 | grep "attn_fwd"; done
echo "== alt 1 CTA/SM"; VST_LIB_PATH=$ALT VST_ATTN_CTAS_PER_SM=1 timeout 300 python scripts/kbench.py attn 2>&1 | grep "attn_fwd"
echo "== exact kernel"; VST_ATTN_STREAM=0 timeout 300 python scripts/kbench.py attn 2>&1 | grep "attn_fwd"
