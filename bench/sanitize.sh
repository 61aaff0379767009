#!/bin/bash
# compute-sanitizer pass over the native kernels (one tool per run: memcheck | racecheck | synccheck | initcheck).
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 200 python bench/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 || { tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 900 compute-sanitizer --tool "$TOOL" --print-limit 20 python bench/sanitize_target.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "exit $?"; grep -c "Invalid\|Race\|hazard\|Uninitialized" gpurun_out/sanitize_$TOOL.log; tail -6 gpurun_out/sanitize_$TOOL.log
