#!/bin/bash
# CTA-pair bring-up: the pair conv tests under a short timeout (a hang must not hold the box)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -k "cta_pair" 2>&1 | tail -30 | tee gpurun_out/pair_r2x.txt
