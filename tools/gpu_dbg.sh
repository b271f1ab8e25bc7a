#!/bin/bash
# Attribution experiments (AQUA_DBG: 1 = no TMA stores, 2 = no epilogue math/staging, 3 = both).
TAG=${1:-x}
mkdir -p gpurun_out
{
for shp in "64 320 320 16 48 3 1 --mblk 4" "64 160 160 48 48 1 1 --mblk 2" "64 320 320 48 96 3 2" "64 160 160 96 96 1 1 --mblk 2" "64 80 80 192 24 1 1"; do
  for dbg in 0 1 2 3; do
    echo "=== $shp dbg=$dbg"
    AQUA_DBG=$dbg timeout 120 python tools/one_op.py $shp 2>&1 | tail -1 | cut -c1-230
  done
done
} > gpurun_out/dbg_$TAG.txt 2>&1
cat gpurun_out/dbg_$TAG.txt
