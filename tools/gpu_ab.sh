#!/bin/bash
# A/B of library builds on single layers: usage tools/gpu_ab.sh <tag> lib1.so lib2.so ...   ("-" = the in-tree build)
TAG=${1:-ab}; shift
mkdir -p gpurun_out
{
for shp in "64 40 40 192 192 3 1 --mblk 1 --mode 1" "64 80 80 96 96 3 1 --mblk 2 --mode 1" "64 160 160 48 48 3 1 --mblk 2 --mode 1" "64 80 80 192 192 1 1 --mblk 1"; do
  for rep in 1 2; do
    for lib in "$@"; do
      if [ "$lib" = "-" ]; then unset AQUA_LIB_PATH; else export AQUA_LIB_PATH=$PWD/$lib; fi
      echo "$lib: $(timeout 120 python tools/one_op.py $shp --iters 30 2>&1 | tail -1 | cut -c1-150)"
    done
  done
done
} 2>&1 | tee gpurun_out/ab_$TAG.txt
