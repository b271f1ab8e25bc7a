#!/bin/bash
# Timeline traces of representative conv shapes.  usage: tools/gpu_trace.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out
{
for shp in "64 320 320 16 48 3 1" "64 160 160 48 48 3 1" "64 80 80 96 96 3 1" "64 40 40 192 192 3 1" "64 20 20 384 384 3 1" \
           "64 320 320 48 96 3 2" "64 160 160 96 192 3 2" "64 40 40 384 768 3 2" \
           "64 160 160 96 96 1 1" "64 160 160 48 48 1 1" "64 80 80 192 192 1 1" "64 40 40 384 384 1 1" "64 40 40 192 192 1 1" "64 20 20 768 768 1 1" "64 20 20 1536 768 1 1"; do
  echo "=== $shp"
  timeout 120 python tools/trace_op.py $shp --tiles 6 2>&1 | tail -40
done
} > gpurun_out/trace_$TAG.txt 2>&1
tail -5 gpurun_out/trace_$TAG.txt
