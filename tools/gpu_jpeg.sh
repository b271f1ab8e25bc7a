#!/bin/bash
# JPEG ingest on the box: parity tests, bench, ncu launch list + full capture of the JPEG kernels.  usage: tools/gpu_jpeg.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_jpeg.py -m gpu -x -q 2>&1 | tail -5
timeout 400 python tools/jpeg_bench.py --e2e 8 > gpurun_out/jpeg_bench_$TAG.json 2> gpurun_out/jpeg_bench_$TAG.err; echo "bench exit=$?"; cat gpurun_out/jpeg_bench_$TAG.json; tail -3 gpurun_out/jpeg_bench_$TAG.err
if [ -z "$NO_NCU" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:jpeg_ -c 60 --csv --log-file gpurun_out/jpeg_launches_$TAG.csv \
    python tools/jpeg_bench.py --iters 2 > gpurun_out/jpeg_ncu_launch_$TAG.log 2>&1; echo "ncu launches exit=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:jpeg_ -s 22 -c 11 -f -o gpurun_out/prof_jpeg_$TAG \
    python tools/jpeg_bench.py --iters 2 > gpurun_out/jpeg_ncu_full_$TAG.log 2>&1; echo "ncu full exit=$?"
fi
