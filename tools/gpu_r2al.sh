#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "raster or windows or blank or stats" > gpurun_out/pytest_win_r2al.log 2>&1; echo "pytest exit=$? $(tail -1 gpurun_out/pytest_win_r2al.log)"; grep -E "Error|FAILED" gpurun_out/pytest_win_r2al.log | cut -c1-300 | head
