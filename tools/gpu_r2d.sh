#!/bin/bash
mkdir -p gpurun_out
for pdl in 1 0; do AQUA_PDL=$pdl timeout 600 python tools/race_probe.py --k 1 2 3 >> gpurun_out/race_r2d.txt 2>> gpurun_out/race_r2d.err; echo "exit=$?"; done
AQUA_NMS_MASK=kernel timeout 600 python tools/race_probe.py --k 1 2 >> gpurun_out/race_r2d.txt 2>> gpurun_out/race_r2d.err; echo "exit=$?"
cat gpurun_out/race_r2d.txt; tail -5 gpurun_out/race_r2d.err
