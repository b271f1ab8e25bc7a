#!/bin/bash
# round 2, visit B: launch-overhead probe, tiling table, parity at the benchmarked configuration, bench
mkdir -p gpurun_out
timeout 600 python tools/launch_overhead.py > gpurun_out/launch_overhead_r2b.txt 2> gpurun_out/launch_overhead_r2b.err; echo "overhead exit=$?"; tail -3 gpurun_out/launch_overhead_r2b.txt
timeout 900 python tools/make_tune_table.py gpurun_out/b200_tune.txt > gpurun_out/tune_build_r2b.log 2>&1; echo "tune exit=$?"; tail -2 gpurun_out/tune_build_r2b.log
mkdir -p aquaculture_b200/tuned && cp gpurun_out/b200_tune.txt aquaculture_b200/tuned/b200.txt
timeout 1200 python -m pytest tests/test_parity_b64.py -m gpu -q -x --tb=short > gpurun_out/parity_b64_r2b.log 2>&1; echo "parity exit=$? $(tail -1 gpurun_out/parity_b64_r2b.log)"
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x --tb=short -k "staged or e2e or cli" > gpurun_out/staged_r2b.log 2>&1; echo "staged exit=$? $(tail -1 gpurun_out/staged_r2b.log)"
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_r2b.json 2> gpurun_out/bench_r2b.err; echo "bench exit=$?"; cat gpurun_out/bench_r2b.json
