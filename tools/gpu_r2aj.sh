#!/bin/bash
mkdir -p gpurun_out
AQUA_TUNE_LOG=0 timeout 300 python tools/jpeg_breakdown.py > gpurun_out/jpeg_breakdown_r2aj.txt 2>&1; cat gpurun_out/jpeg_breakdown_r2aj.txt | tail -15
