#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/ab_env.py AQUA_CHAIN=1 - AQUA_CHAIN=0 > gpurun_out/ab_chain_r2n.txt 2> gpurun_out/ab_chain_r2n.err; cat gpurun_out/ab_chain_r2n.txt
timeout 600 python tools/ab_env.py AQUA_CHAIN=0 AQUA_CHAIN=1 >> gpurun_out/ab_chain_r2n.txt 2>> gpurun_out/ab_chain_r2n.err; tail -2 gpurun_out/ab_chain_r2n.txt
AQUA_CHAIN=1 timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-jpeg --no-extra-configs --parity-tiles 0 > gpurun_out/bench_chain1.json 2>/dev/null
AQUA_CHAIN=0 timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-jpeg --no-extra-configs --parity-tiles 0 > gpurun_out/bench_chain0.json 2>/dev/null
AQUA_CHAIN=1 timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-jpeg --no-extra-configs --parity-tiles 0 > gpurun_out/bench_chain1b.json 2>/dev/null
AQUA_CHAIN=0 timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-jpeg --no-extra-configs --parity-tiles 0 > gpurun_out/bench_chain0b.json 2>/dev/null
for f in chain1 chain0 chain1b chain0b; do python -c "
import json; d=json.load(open('gpurun_out/bench_$f.json')); print('$f', d['value'], d['ms_per_step'], d['roofline']['ms_per_conv_stack'], d['clocks']['sm_mhz'])"; done
