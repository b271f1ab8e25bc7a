#!/bin/bash
mkdir -p gpurun_out
python tools/weights_hash.py
timeout 900 python tools/repro_probe.py --iters 6000 2>> gpurun_out/soak.err | grep -v "^[0-9]* conv\|^[0-9]* detect\|^[0-9]* sppf" | cut -c1-200
timeout 600 python tools/race_probe.py 2>> gpurun_out/soak.err | tail -3 | cut -c1-200
