#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/repro_r2f.txt gpurun_out/race_r2f.txt
echo "=== default (per-set barrier ids)" >> gpurun_out/repro_r2f.txt
timeout 900 python tools/repro_probe.py --iters 2000 2>> gpurun_out/repro_r2f.err | grep -v "^[0-9]* conv\|^[0-9]* detect\|^[0-9]* sppf" | cut -c1-300 >> gpurun_out/repro_r2f.txt; echo "exit=$?"
cat gpurun_out/repro_r2f.txt
for pdl in 1 0; do AQUA_PDL=$pdl timeout 600 python tools/race_probe.py --batches 160 --k 1 2 3 >> gpurun_out/race_r2f.txt 2>> gpurun_out/race_r2f.err; echo "exit=$?"; done
AQUA_NMS_MASK=kernel timeout 600 python tools/race_probe.py --batches 160 --k 1 2 >> gpurun_out/race_r2f.txt 2>> gpurun_out/race_r2f.err; echo "exit=$?"
cat gpurun_out/race_r2f.txt; tail -5 gpurun_out/race_r2f.err
