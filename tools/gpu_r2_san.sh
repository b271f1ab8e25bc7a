#!/bin/bash
# one compute-sanitizer tool per gpurun call: tools/gpu_r2_san.sh memcheck|racecheck|synccheck
TOOL=${1:-memcheck}
mkdir -p gpurun_out
export AQUA_TUNE_LOG=0
timeout 600 python tools/sanitize_prog.py > gpurun_out/san_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/san_plain.log; exit 1; }
tail -8 gpurun_out/san_plain.log
timeout 3000 compute-sanitizer --tool $TOOL --print-limit 40 --log-file gpurun_out/san_$TOOL.log python tools/sanitize_prog.py > gpurun_out/san_${TOOL}_stdout.log 2>&1
echo "sanitizer $TOOL exit=$?"
tail -8 gpurun_out/san_${TOOL}_stdout.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error|hazard" gpurun_out/san_$TOOL.log | sort | uniq -c | sort -rn | head -20
