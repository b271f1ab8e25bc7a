#!/bin/bash
TAG=${1:-s16b}
mkdir -p gpurun_out
run() { timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --tb=short -k "$2" > gpurun_out/selftest_$1.log 2>&1; echo "$1 exit=$? $(tail -1 gpurun_out/selftest_$1.log)"; }
run cluster "cluster"
tail -30 gpurun_out/selftest_cluster.log | cut -c1-300
{
for shp in "64 160 160 96 192 3 2" "64 80 80 192 384 3 2" "64 40 40 384 768 3 2" "64 80 80 192 192 3 2" "64 40 40 384 384 1 1" "64 40 40 768 384 1 1" "64 20 20 768 768 1 1" "64 20 20 1536 768 1 1" "64 40 40 192 192 3 1" "64 20 20 384 384 3 1" "64 80 80 96 96 3 1"; do
  for cl in 1 2 4; do
    timeout 120 python tools/one_op.py $shp --cluster $cl 2>&1 | tail -1 | cut -c1-220
  done
done
} 2>&1 | tee gpurun_out/cluster_sweep_$TAG.txt
