#!/bin/bash
# Parity suite + per-op table + bench (no ncu).  usage: tools/gpu_quick.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out
bash tools/gpu_selftest.sh 2>&1 | tee gpurun_out/selftest_$TAG.txt
timeout 300 python tools/profile_ops.py --out gpurun_out/ops_$TAG.txt > /dev/null 2> gpurun_out/ops_$TAG.err; tail -1 gpurun_out/ops_$TAG.txt
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit=$?"; cut -c1-400 gpurun_out/bench_$TAG.json
