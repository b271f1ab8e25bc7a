#!/bin/bash
# Two-set epilogue: parity suite with the organisation pinned, then one-op timings 1 vs 2 sets.
export AQUA_EPI_SETS=2
bash tools/gpu_selftest.sh 2>&1 | grep -v "nms exit=0\|prepost exit=0"
unset AQUA_EPI_SETS
{
for shp in "64 320 320 16 48 3 1" "64 160 160 48 48 1 1" "64 160 160 96 96 1 1" "64 160 160 48 48 3 1" "64 320 320 48 96 3 2" "64 80 80 192 192 1 1" "64 80 80 96 96 3 1" "64 40 40 192 192 3 1" "64 80 80 96 96 1 1"; do
  for sets in 1 2; do
    echo "=== $shp sets=$sets"
    AQUA_EPI_SETS=$sets timeout 120 python tools/one_op.py $shp 2>&1 | tail -1 | cut -c1-200
  done
done
} > gpurun_out/sets_$1.txt 2>&1
cat gpurun_out/sets_$1.txt
