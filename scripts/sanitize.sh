#!/bin/bash
# compute-sanitizer over one small invocation of every kernel family (scripts/sanitize_cases.py).
# ONE tool per gpurun call (B200_PROFILING.md: several tools in one call have wedged the GPU):
#   gpurun -- 'bash scripts/sanitize.sh memcheck'     (then racecheck, synccheck in separate calls)
# The program is first run plain; the sanitizer only runs if that exits 0.  Log -> gpurun_out/sanitize_<tool>.log
TOOL=${1:-memcheck}
mkdir -p gpurun_out
python scripts/sanitize_cases.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 50 python scripts/sanitize_cases.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error:|hazard" gpurun_out/sanitize_$TOOL.log | head -20
tail -n 5 gpurun_out/sanitize_$TOOL.log
