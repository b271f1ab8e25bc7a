#!/bin/bash
mkdir -p gpurun_out
AQUA_TUNE_LOG=0 timeout 300 python tools/profile_ops.py --out gpurun_out/ops_r2ac.txt > /dev/null 2> gpurun_out/ops_r2ac.err
AQUA_TUNE_LOG=0 timeout 300 python tools/profile_ops.py --insitu --out gpurun_out/ops_insitu_r2ac.txt > /dev/null 2>> gpurun_out/ops_r2ac.err
tail -1 gpurun_out/ops_r2ac.txt; tail -1 gpurun_out/ops_insitu_r2ac.txt
