#!/bin/bash
TAG=${1:-s16c}
mkdir -p gpurun_out
run() { timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --tb=short -k "$2" > gpurun_out/selftest_$1.log 2>&1; echo "$1 exit=$? $(tail -1 gpurun_out/selftest_$1.log)"; }
run s2variants "s2_variants"
tail -30 gpurun_out/selftest_s2variants.log | cut -c1-300
run conv3x3s2 "conv3x3_s2"
{
for shp in "64 320 320 48 96 3 2" "64 160 160 96 192 3 2" "64 80 80 192 384 3 2" "64 40 40 384 768 3 2" "64 80 80 192 192 3 2" "64 40 40 384 384 3 2"; do
  for md in "--mode 2" "--mode 1 --mblk 1" "--mode 1 --mblk 2" "--mode 1 --mblk 4"; do
    timeout 120 python tools/one_op.py $shp $md 2>&1 | tail -1 | cut -c1-220
  done
done
} 2>&1 | tee gpurun_out/halo2_sweep_$TAG.txt
