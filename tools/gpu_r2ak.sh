#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "nms" 2>&1 | tail -2
AQUA_TUNE_LOG=0 timeout 900 python tools/cli_jpeg_bench.py --tiles 6144 > gpurun_out/cli_jpeg_r2ak.json 2> gpurun_out/cli_jpeg_r2ak.err; echo "cli bench exit=$?"; cut -c1-1500 gpurun_out/cli_jpeg_r2ak.json; tail -3 gpurun_out/cli_jpeg_r2ak.err
