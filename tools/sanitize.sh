#!/bin/bash
# compute-sanitizer over one small pass of every kernel (tools/sanitize_case.py).  ONE tool per gpurun call
# (B200_PROFILING.md: several tools in one call have left a GPU unusable):
#     gpurun -- 'bash tools/sanitize.sh memcheck'      (then racecheck, initcheck, synccheck in later calls)
# STATUS on this pool (round 2): the plain run passes (116 launches, results checked against the CPU oracle), but
# `compute-sanitizer` answers "closed on this pool and stays closed" (exit 86, gpurun_out/sanitize_memcheck.log), so no
# sanitizer summary could be committed; the script is kept for a pool where the tool is open.
# The plain run must pass first; the summary line of the tool goes to gpurun_out/sanitize_<tool>.log.
TOOL=${1:-memcheck}
mkdir -p gpurun_out
export FSB_PDL=0   # programmatic dependent launch overlaps kernels by design; the sanitizer serialises launches anyway
timeout 300 python tools/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
tail -1 gpurun_out/sanitize_plain.log
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python tools/sanitize_case.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize case ok|Error|error" gpurun_out/sanitize_$TOOL.log | head -20
