#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu_r2ag.log 2>&1; echo "pytest -m gpu exit=$? $(tail -1 gpurun_out/pytest_gpu_r2ag.log)"; grep -E "Error|FAILED" gpurun_out/pytest_gpu_r2ag.log | cut -c1-400 | head
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_r2ag.json 2> gpurun_out/bench_r2ag.err; echo "bench exit=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_r2ag.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['ms_per_conv_stack'], d['parity']['ok'], d['aux_kernels']['nms_kernel'], d['extra_configs']['nms_stress'], d['stage_ms_last_batch'], d['clocks'])"
