#!/bin/bash
# vst_gemm tile width chosen inside the instantiation (N = 320 -> 2 x 160, 640 -> 3 x 224): parity, bwd microbench, cfg4
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_backward_gpu.py tests/test_parity_full_gpu.py -m gpu -q -x > gpurun_out/tests_r.log 2>&1; echo "tests rc=$?"; tail -n 3 gpurun_out/tests_r.log | cut -c1-250
python scripts/kbench.py bwd > gpurun_out/kbench_bwd_r.log 2>&1; grep -E "vst_gemm|cuBLAS" gpurun_out/kbench_bwd_r.log
VST_KERNEL_TABLE=gpurun_out/cfg4_kernels_r.tsv python bench.py --workload cfg4 --steps 3 > gpurun_out/bench_cfg4_r.json 2> gpurun_out/bench_cfg4_r.err; echo "cfg4 rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_cfg4_r.json').read().strip().splitlines()[-1]); print('cfg4 ms', d['ms_per_step'], d['clocks'])"
grep -E "gemm_t_kernel|gemm_reduce" gpurun_out/cfg4_kernels_r.tsv | cut -c1-60,150-220
