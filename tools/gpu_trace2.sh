#!/bin/bash
TAG=${1:-x}; shift
mkdir -p gpurun_out
{
for shp in "$@"; do
  echo "=== $shp"
  timeout 120 python tools/trace_op.py $shp --cta 0,73 --tiles 5 2>&1 | tail -22
done
} > gpurun_out/trace_$TAG.txt 2>&1
