#!/bin/bash
# GEGLU epilogue A/B: one-MUFU erf (main build) vs rcp+ex2 erf (alt build)
mkdir -p gpurun_out
ALT=$PWD/video_style_transfer_b200/libvst_b200_alt.so
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_modules_gpu.py -m gpu -q -x -k "geglu or motion or block" > gpurun_out/tests_m.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/tests_m.log
echo "== main: one MUFU"; python scripts/kbench.py geglu 2>&1 | grep GEGLU
echo "== alt: rcp + ex2"; VST_LIB_PATH=$ALT python scripts/kbench.py geglu 2>&1 | grep GEGLU
echo "== main again"; python scripts/kbench.py geglu 2>&1 | grep GEGLU
