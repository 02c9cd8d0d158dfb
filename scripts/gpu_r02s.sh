#!/bin/bash
# CTA-pair vst_gemm (gemm_t2_kernel): parity, bwd microbench pair on / off, cfg4
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_backward_gpu.py tests/test_parity_full_gpu.py -m gpu -q -x > gpurun_out/tests_s.log 2>&1; echo "tests rc=$?"; tail -n 5 gpurun_out/tests_s.log | cut -c1-300
echo "== pair"; python scripts/kbench.py bwd 2>&1 | grep -E "vst_gemm|cuBLAS" | tee gpurun_out/kbench_bwd_s.log
echo "== single"; VST_GEMM_T_PAIR=0 python scripts/kbench.py bwd 2>&1 | grep -E "vst_gemm" | tee gpurun_out/kbench_bwd_s_single.log
VST_KERNEL_TABLE=gpurun_out/cfg4_kernels_s.tsv python bench.py --workload cfg4 --steps 3 > gpurun_out/bench_cfg4_s.json 2> gpurun_out/bench_cfg4_s.err; echo "cfg4 rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench_cfg4_s.json').read().strip().splitlines()[-1]); print('cfg4 ms', d['ms_per_step'], d['clocks'])"
grep -E "gemm_t|gemm_reduce|# one" gpurun_out/cfg4_kernels_s.tsv | awk -F'\t' '{print substr($1,1,60), $2, $3, $4}'
