#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -q -m gpu -k "nms or parity or detect" > gpurun_out/pytest_nms_r2ae.log 2>&1; echo "pytest exit=$? $(tail -1 gpurun_out/pytest_nms_r2ae.log)"; grep -E "Error|FAILED" gpurun_out/pytest_nms_r2ae.log | cut -c1-300 | head
AQUA_TUNE_LOG=0 timeout 300 python tools/nms_stress.py > gpurun_out/nms_stress_r2ae.txt 2>&1
AQUA_NMS_CLK=1 AQUA_TUNE_LOG=0 timeout 300 python tools/nms_stress.py --iters 3 2>&1 | tail -2 >> gpurun_out/nms_stress_r2ae.txt
cat gpurun_out/nms_stress_r2ae.txt
