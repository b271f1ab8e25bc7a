#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_chain.py -m gpu -q -x --tb=short > gpurun_out/chain_r2k.log 2>&1; echo "chain tests exit=$? $(tail -1 gpurun_out/chain_r2k.log)"; tail -25 gpurun_out/chain_r2k.log | head -40
timeout 600 python tools/ab_env.py - AQUA_CHAIN=1 > gpurun_out/ab_chain_r2k.txt 2> gpurun_out/ab_chain_r2k.err; cat gpurun_out/ab_chain_r2k.txt; tail -3 gpurun_out/ab_chain_r2k.err
