#!/bin/bash
mkdir -p gpurun_out
AQUA_TUNE_LOG=0 timeout 900 python tools/ab_env.py - AQUA_DET_LATE=1 > gpurun_out/ab_det_r2w.txt 2> gpurun_out/ab_det_r2w.err; cat gpurun_out/ab_det_r2w.txt | cut -c1-170; tail -2 gpurun_out/ab_det_r2w.err
