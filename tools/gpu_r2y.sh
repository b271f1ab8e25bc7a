#!/bin/bash
# CTA pairs as autotune candidates: tune the batch-64 plan with them, then A/B against the shipped single-CTA tilings
mkdir -p gpurun_out
TUNE_CHAINS="feed,c3" AQUA_TUNE_LOG=2 timeout 600 python tools/make_tune_table.py gpurun_out/pair_tune.txt 64:640 > gpurun_out/pair_tune_r2y.log 2>&1; echo "tune exit=$?"; tail -2 gpurun_out/pair_tune_r2y.log
cat aquaculture_b200/tuned/b200.txt gpurun_out/pair_tune.txt > gpurun_out/combined_tune.txt
AQUA_TUNE_TABLE=gpurun_out/combined_tune.txt AQUA_TUNE_LOG=0 timeout 900 python tools/ab_env.py AQUA_PAIR=0 - > gpurun_out/ab_pair_r2y.txt 2> gpurun_out/ab_pair_r2y.err; cat gpurun_out/ab_pair_r2y.txt | cut -c1-170; tail -2 gpurun_out/ab_pair_r2y.err
