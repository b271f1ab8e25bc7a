#!/bin/bash
mkdir -p gpurun_out
AQUA_NMS_CLK=1 timeout 300 python tools/forward_only.py --warm 0 --passes 0 > gpurun_out/nms_clk.txt 2>&1; grep "nms phases" gpurun_out/nms_clk.txt | head -3
