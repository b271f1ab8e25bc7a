#!/bin/bash
# round 2, visit A: parity suite with the staged plan, then the stage sweep
mkdir -p gpurun_out
bash tools/gpu_selftest.sh 2>&1 | tee gpurun_out/selftest_r2a.txt
export AQUA_TUNE_CACHE=gpurun_out/tune_r2a.txt
timeout 1500 python tools/stage_sweep.py 0 "a=4x2,b=16x2,c=16x2" "a=4x1,b=16x1,c=16x1" "a=8x2,b=16x2,c=16x2" "a=8x1,b=16x1,c=16x1" \
  "a=4x2,b=0,c=0" "a=0,b=16x2,c=0" "a=0,b=0,c=16x2" "a=2x2,b=8x2,c=8x2" "a=4x3,b=16x3,c=16x3" "a=8x2,b=32x2,c=32x2" "a=16x2,b=32x2,c=32x2" \
  > gpurun_out/stage_sweep_r2a.txt 2> gpurun_out/stage_sweep_r2a.err
echo "sweep exit=$?"; cat gpurun_out/stage_sweep_r2a.txt; tail -5 gpurun_out/stage_sweep_r2a.err
