#!/bin/bash
# session-16 GPU visit: parity suite, pipelined bench, launch list of the real step, probes, traces
TAG=${1:-s16}
mkdir -p gpurun_out
bash tools/gpu_selftest.sh 2>&1 | tee gpurun_out/selftest_$TAG.txt
export AQUA_TUNE_CACHE=gpurun_out/tune_$TAG.txt
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit=$?"; cut -c1-1500 gpurun_out/bench_$TAG.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch_$TAG.log 2>&1
echo "ncu launches exit=$?"
unset AQUA_TUNE_CACHE
./tools/umma_probe b > gpurun_out/probe_t11_$TAG.txt 2>&1; cat gpurun_out/probe_t11_$TAG.txt
for a in 0 2; do timeout 60 python tools/one_op.py 64 320 320 16 48 3 1 --act $a 2>&1 | tail -1 | cut -c1-200; done
for a in 0 2; do timeout 60 python tools/one_op.py 64 160 160 48 48 1 1 --act $a 2>&1 | tail -1 | cut -c1-200; done
for a in 0 2; do timeout 60 python tools/one_op.py 64 80 80 96 96 1 1 --act $a 2>&1 | tail -1 | cut -c1-200; done
bash tools/gpu_trace2.sh $TAG "64 320 320 16 48 3 1" "64 320 320 48 96 3 2" "64 160 160 96 192 3 2" "64 80 80 192 384 3 2" "64 40 40 384 768 3 2" "64 40 40 384 384 1 1" "64 40 40 768 384 1 1" "64 20 20 768 768 1 1" "64 80 80 96 96 1 1"
