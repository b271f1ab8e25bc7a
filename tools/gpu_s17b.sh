#!/bin/bash
TAG=${1:-s17b}
mkdir -p gpurun_out
run() { timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --tb=short -k "$2" > gpurun_out/selftest_$1.log 2>&1; echo "$1 exit=$? $(tail -1 gpurun_out/selftest_$1.log)"; }
run folded "folded_upsample"
grep -n "Error\|FAILED" gpurun_out/selftest_folded.log | head -5 | cut -c1-300
run s2variants "s2_variants"
grep -n "Error\|FAILED" gpurun_out/selftest_s2variants.log | head -5 | cut -c1-300
run e2e "e2e or label_files"
export AQUA_TUNE_CACHE=gpurun_out/tune_$TAG.txt
timeout 300 python tools/profile_ops.py --out gpurun_out/ops_fold_$TAG.txt > /dev/null 2> gpurun_out/ops_$TAG.err; tail -1 gpurun_out/ops_fold_$TAG.txt
AQUA_FOLD_UPSAMPLE=0 timeout 300 python tools/profile_ops.py --out gpurun_out/ops_nofold_$TAG.txt > /dev/null 2>> gpurun_out/ops_$TAG.err; tail -1 gpurun_out/ops_nofold_$TAG.txt
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit=$?"; cut -c1-300 gpurun_out/bench_$TAG.json; grep -o '"roofline.*' gpurun_out/bench_$TAG.json | cut -c1-400
