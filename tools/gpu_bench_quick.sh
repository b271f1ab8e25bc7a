#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_quick.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['traffic'], d['extra_configs']['img1280'], d['extra_configs']['nms_stress']['nms_ms_per_64_images'])"; tail -2 gpurun_out/bench_quick.err
