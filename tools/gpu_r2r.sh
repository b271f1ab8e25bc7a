#!/bin/bash
mkdir -p gpurun_out
AQUA_TUNE_LOG=0 timeout 1200 python tools/ab_env.py - AQUA_STAGES=a=32x1 AQUA_STAGES=a=16x1 AQUA_STAGES=b=32x1 AQUA_STAGES=a=32x1+AQUA_CHAIN=c3 AQUA_STAGES=a=32x1,b=32x1,c=32x1 > gpurun_out/ab_stage_r2r.txt 2> gpurun_out/ab_stage_r2r.err; cat gpurun_out/ab_stage_r2r.txt | cut -c1-170; tail -3 gpurun_out/ab_stage_r2r.err
