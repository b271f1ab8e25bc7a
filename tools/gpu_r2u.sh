#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-jpeg --parity-tiles 0 --no-extra-configs > gpurun_out/bench_r2u.json 2> gpurun_out/bench_r2u.err; python -c "
import json; d=json.load(open('gpurun_out/bench_r2u.json')); print(d['value'], d['ms_per_step'], d['roofline']['ms_per_conv_stack'], d['stage_ms_last_batch'], d['aux_kernels']['nms_kernel']['us'], d['aux_kernels']['preprocess_kernel']['us'])"
timeout 300 python tools/forward_only.py > gpurun_out/fwd_plain2.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'nms|preprocess|sppf' -c 12 --csv --log-file gpurun_out/nms_launches.csv python tools/forward_only.py > gpurun_out/ncu_nms.log 2>&1
grep -E "nms|preprocess|sppf" gpurun_out/nms_launches.csv | cut -d, -f5,15 | head -12
