#!/bin/bash
mkdir -p gpurun_out
for m in 1; do
AQUA_NMS_CLK=$m AQUA_TUNE_LOG=0 timeout 300 python tools/nms_stress.py --iters 2 2>&1 | grep "nms " | tail -2
done > gpurun_out/nms_stress_r2af.txt
cat gpurun_out/nms_stress_r2af.txt
