#!/bin/bash
mkdir -p gpurun_out
{
echo "== stem 3x3 16->48 @320"; AQUA_TUNE_LOG=0 timeout 120 python tools/trace_op.py 64 320 320 16 48 3 1 --cta 0,73 2>&1 | tail -25
echo "== 1x1 48->48 @160 (ct 96)"; AQUA_TUNE_LOG=0 timeout 120 python tools/trace_op.py 64 160 160 48 48 1 1 --cta 0,73 2>&1 | tail -25
echo "== 3x3 48->48 res @160"; AQUA_TUNE_LOG=0 timeout 120 python tools/trace_op.py 64 160 160 48 48 3 1 --res --cta 0,73 2>&1 | tail -25
echo "== 3x3 96->96 res @80"; AQUA_TUNE_LOG=0 timeout 120 python tools/trace_op.py 64 80 80 96 96 3 1 --res --cta 0,73 2>&1 | tail -25
} > gpurun_out/trace_r2ah.txt
wc -l gpurun_out/trace_r2ah.txt
