#!/bin/bash
mkdir -p gpurun_out
AQUA_CHAIN=1 timeout 300 python tools/profile_ops.py --out gpurun_out/ops_chain_r2l.txt > /dev/null 2> gpurun_out/ops_chain_r2l.err
grep -n "+1x1\|TOTAL" gpurun_out/ops_chain_r2l.txt | cut -c1-260
timeout 300 python tools/profile_ops.py --out gpurun_out/ops_plain_r2l.txt > /dev/null 2>> gpurun_out/ops_chain_r2l.err
sed -n 3,4p gpurun_out/ops_plain_r2l.txt | cut -c1-200; sed -n 10,11p gpurun_out/ops_plain_r2l.txt | cut -c1-200; sed -n 53,54p gpurun_out/ops_plain_r2l.txt | cut -c1-200; tail -1 gpurun_out/ops_plain_r2l.txt
