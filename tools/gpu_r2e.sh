#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/repro_r2e.txt
for v in "A=1" "AQUA_ES2_EVEN_STG=1" "AQUA_PDL_DET_OFF=1"; do
  echo "=== $v" >> gpurun_out/repro_r2e.txt
  env $v timeout 900 python tools/repro_probe.py --iters 600 2>> gpurun_out/repro_r2e.err | grep -v "^[0-9]* conv\|^[0-9]* detect\|^[0-9]* sppf" | cut -c1-400 >> gpurun_out/repro_r2e.txt; echo "exit=$?"
done
cat gpurun_out/repro_r2e.txt; tail -3 gpurun_out/repro_r2e.err
