#!/bin/bash
mkdir -p gpurun_out
rm -f gpurun_out/ab_sleep_r2q.txt
for r in 1 2 3; do
  for lib in aquaculture_b200/libaqua_b200.so aquaculture_b200/libaqua_b200_sleep.so; do
    echo "== $lib" >> gpurun_out/ab_sleep_r2q.txt
    AQUA_LIB_PATH=$PWD/$lib AB_ROUNDS=5 AQUA_TUNE_LOG=0 timeout 300 python tools/ab_env.py - >> gpurun_out/ab_sleep_r2q.txt 2>> gpurun_out/ab_sleep_r2q.err
  done
done
cat gpurun_out/ab_sleep_r2q.txt | cut -c1-150; tail -2 gpurun_out/ab_sleep_r2q.err
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_r2q.log 2>&1; echo "pytest -m gpu exit=$? $(tail -1 gpurun_out/pytest_gpu_r2q.log)"; grep -E "Error|FAILED" gpurun_out/pytest_gpu_r2q.log | cut -c1-400 | head
