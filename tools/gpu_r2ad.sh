#!/bin/bash
mkdir -p gpurun_out
AQUA_TUNE_LOG=0 timeout 300 python tools/nms_stress.py > gpurun_out/nms_stress_r2ad.txt 2>&1
AQUA_NMS_CLK=1 AQUA_TUNE_LOG=0 timeout 300 python tools/nms_stress.py --iters 3 2>&1 | tail -4 >> gpurun_out/nms_stress_r2ad.txt
cat gpurun_out/nms_stress_r2ad.txt
