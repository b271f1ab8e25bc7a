#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_chain.py -m gpu -q --tb=short > gpurun_out/chain_r2o.log 2>&1; echo "chain tests exit=$? $(tail -1 gpurun_out/chain_r2o.log)"; grep -E "AssertionError|FAILED" gpurun_out/chain_r2o.log | cut -c1-700 | head -12
AQUA_TUNE_LOG=0 timeout 900 python tools/ab_env.py AQUA_CHAIN=0 AQUA_CHAIN=c3 AQUA_CHAIN=feed AQUA_CHAIN=feed,c3 AQUA_SERPENTINE=1 > gpurun_out/ab_chain_r2o.txt 2> gpurun_out/ab_chain_r2o.err; cat gpurun_out/ab_chain_r2o.txt; tail -3 gpurun_out/ab_chain_r2o.err
AQUA_CHAIN=1 AQUA_TUNE_LOG=0 timeout 300 python tools/profile_ops.py --out gpurun_out/ops_chain_r2o.txt > /dev/null 2>> gpurun_out/ab_chain_r2o.err
grep -n "+1x1\|TOTAL" gpurun_out/ops_chain_r2o.txt | cut -c1-250
