#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python tools/make_tune_table.py gpurun_out/b200_tune.txt > gpurun_out/tune_build_r2aa.log 2>&1; echo "tune exit=$?"; tail -2 gpurun_out/tune_build_r2aa.log
cp gpurun_out/b200_tune.txt aquaculture_b200/tuned/b200.txt
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_r2aa.log 2>&1; echo "pytest -m gpu exit=$? $(tail -1 gpurun_out/pytest_gpu_r2aa.log)"; grep -E "Error|FAILED" gpurun_out/pytest_gpu_r2aa.log | cut -c1-400 | head
timeout 300 python tools/repro_probe.py --iters 500 2>> gpurun_out/repro_r2aa.err | grep -v "^[0-9]* conv\|^[0-9]* detect\|^[0-9]* sppf" | cut -c1-300 > gpurun_out/repro_r2aa.txt; cat gpurun_out/repro_r2aa.txt
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_r2aa.json 2> gpurun_out/bench_r2aa.err; echo "bench exit=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_r2aa.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['ms_per_conv_stack'], d['parity']['ok'], d['parity'].get('within_tolerance'), d['tilings'], d['clocks'])"
