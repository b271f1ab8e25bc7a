#!/bin/bash
# Forced tilings: fewer weight bytes per output (2 M blocks per tile, narrower N tiles) vs the autotuned choice.
TAG=${1:-x}
{
run() { echo "=== $*"; timeout 120 python tools/one_op.py $* 2>&1 | tail -1 | cut -c1-210; }
run 64 40 40 192 192 3 1
run 64 40 40 192 192 3 1 --mblk 2
run 64 40 40 192 192 3 1 --mblk 2 --tile-n 96
run 64 40 40 192 192 3 1 --mblk 4 --tile-n 96
run 64 40 40 192 192 3 1 --mblk 4 --tile-n 64
run 64 160 160 96 192 3 2
run 64 160 160 96 192 3 2 --mblk 2
run 64 160 160 96 192 3 2 --mblk 2 --tile-n 96
run 64 20 20 384 384 3 1
run 64 20 20 384 384 3 1 --mblk 2 --tile-n 96
run 64 20 20 384 384 3 1 --mblk 4 --tile-n 64
run 64 40 40 384 768 3 2
run 64 40 40 384 768 3 2 --mblk 2 --tile-n 128
run 64 40 40 384 768 3 2 --mblk 2 --tile-n 96
run 64 20 20 768 768 1 1
run 64 20 20 768 768 1 1 --mblk 2 --tile-n 128
run 64 20 20 1536 768 1 1
run 64 20 20 1536 768 1 1 --mblk 2 --tile-n 128
} > gpurun_out/tn_$TAG.txt 2>&1
cat gpurun_out/tn_$TAG.txt
