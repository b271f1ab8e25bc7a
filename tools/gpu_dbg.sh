#!/bin/bash
# Attribution experiments on single layers with AQUA_DBG (results are WRONG under any non-zero value; timing only):
#   1 = no TMA stores issued, 2 = epilogue skips math + staging, 8 = no TMEM loads, 16 = no proxy fence,
#   32 = no staging hand-off between epilogue and store warp (needs 1), 64 = old per-store scalar path of HALO stores.
# usage: tools/gpu_dbg.sh <tag>   -> gpurun_out/dbg_<tag>.txt   (summary kept in profiles/r1_attribution_stem_v7.txt)
TAG=${1:-x}
mkdir -p gpurun_out
{
for shp in "64 320 320 16 48 3 1 --mblk 4" "64 160 160 48 48 1 1 --mblk 2" "64 320 320 48 96 3 2" "64 160 160 96 96 1 1 --mblk 2"; do
  for dbg in 0 1 2 3 35 59 64; do
    echo "=== $shp dbg=$dbg"
    AQUA_DBG=$dbg timeout 120 python tools/one_op.py $shp 2>&1 | tail -1 | cut -c1-230
  done
done
} > gpurun_out/dbg_$TAG.txt 2>&1
cat gpurun_out/dbg_$TAG.txt
