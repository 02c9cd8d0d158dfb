#!/bin/bash
# the allocation-free entry point: alone in a fresh process, then after the whole suite (pytest -x, as the round's end)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/tests_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -n 6 gpurun_out/tests_noalloc_alone.log | cut -c1-300
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/tests_j.log 2>&1; echo "suite rc=$?"; tail -n 6 gpurun_out/tests_j.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()"; echo "smoke rc=$?"
