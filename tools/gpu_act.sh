#!/bin/bash
# Cost of the activation in the epilogue-bound layers: MUFU probe + one-op timings with act = none / exact / fast.
TAG=${1:-x}
mkdir -p gpurun_out
./tools/mufu_probe > gpurun_out/mufu_$TAG.txt 2>&1
{
for shp in "64 320 320 16 48 3 1" "64 160 160 48 48 1 1" "64 160 160 96 96 1 1" "64 160 160 48 48 3 1" "64 80 80 192 192 1 1"; do
  for act in 0 1 2; do
    echo "=== $shp act=$act"
    timeout 120 python tools/one_op.py $shp --act $act 2>&1 | tail -2
  done
done
} > gpurun_out/act_$TAG.txt 2>&1
cat gpurun_out/mufu_$TAG.txt gpurun_out/act_$TAG.txt
