#!/bin/bash
TAG=${1:-x}
{
for dbg in 0 3 11 19 35 43 59; do
  echo "=== stem dbg=$dbg"
  AQUA_DBG=$dbg timeout 120 python tools/one_op.py 64 320 320 16 48 3 1 --mblk 4 2>&1 | tail -1 | cut -c1-60
done
for dbg in 0 3 35 59; do
  echo "=== 1x1 48 dbg=$dbg"
  AQUA_DBG=$dbg timeout 120 python tools/one_op.py 64 160 160 48 48 1 1 --mblk 2 2>&1 | tail -1 | cut -c1-60
done
} > gpurun_out/dbg2_$TAG.txt 2>&1
cat gpurun_out/dbg2_$TAG.txt
