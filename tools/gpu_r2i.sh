#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/jpeg_breakdown.py > gpurun_out/jpeg_breakdown_r2i.txt 2> gpurun_out/jpeg_breakdown_r2i.err; echo "exit=$?"; cat gpurun_out/jpeg_breakdown_r2i.txt; tail -3 gpurun_out/jpeg_breakdown_r2i.err

