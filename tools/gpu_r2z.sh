#!/bin/bash
mkdir -p gpurun_out
cat aquaculture_b200/tuned/b200.txt aquaculture_b200/tuned/pair_try.txt > /tmp/combined_tune.txt
AQUA_TUNE_TABLE=/tmp/combined_tune.txt AQUA_TUNE_LOG=1 timeout 900 python tools/ab_env.py AQUA_PAIR=0 - > gpurun_out/ab_pair_r2z.txt 2> gpurun_out/ab_pair_r2z.err; cat gpurun_out/ab_pair_r2z.txt | cut -c1-170; tail -2 gpurun_out/ab_pair_r2z.err
