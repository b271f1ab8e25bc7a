#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/ab_env.py AQUA_SERPENTINE=0 - AQUA_SERPENTINE=0 - > gpurun_out/ab_serp_r2g.txt 2> gpurun_out/ab_serp_r2g.err; echo "exit=$?"; cat gpurun_out/ab_serp_r2g.txt; tail -3 gpurun_out/ab_serp_r2g.err
AB_BATCH=16 AB_IMG=1280 timeout 900 python tools/ab_env.py AQUA_SERPENTINE=0 - >> gpurun_out/ab_serp_r2g.txt 2>> gpurun_out/ab_serp_r2g.err; tail -2 gpurun_out/ab_serp_r2g.txt
timeout 600 python tools/repro_probe.py --iters 300 2>> gpurun_out/repro_r2g.err | grep -v "^[0-9]* conv\|^[0-9]* detect\|^[0-9]* sppf" | cut -c1-300 > gpurun_out/repro_r2g.txt; cat gpurun_out/repro_r2g.txt
timeout 900 python -m pytest tests/test_parity_b64.py -m gpu -q --tb=short -x > gpurun_out/parity_r2g.log 2>&1; echo "parity exit=$? $(tail -1 gpurun_out/parity_r2g.log)"
