#!/bin/bash
# DRAM traffic per launch of one bench step (fills roofline.traffic): ncu with the two dram counters, tilings replayed.
# usage: tools/gpu_dram_pass.sh <tag>   (after tools/gpu_round.sh <tag>, which leaves gpurun_out/tune_<tag>.txt)
TAG=${1:-x}
export AQUA_TUNE_CACHE=gpurun_out/tune_$TAG.txt
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 500 --csv \
    --log-file gpurun_out/dram_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_dram_$TAG.log 2>&1
echo "ncu dram exit=$?"
