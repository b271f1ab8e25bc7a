#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu_r2ab.log 2>&1; echo "pytest -m gpu exit=$? $(tail -1 gpurun_out/pytest_gpu_r2ab.log)"; grep -E "Error|FAILED" gpurun_out/pytest_gpu_r2ab.log | cut -c1-400 | head
timeout 600 python tools/repro_probe.py --iters 2000 2>> gpurun_out/repro_r2ab.err | grep -v "^[0-9]* conv\|^[0-9]* detect\|^[0-9]* sppf" | cut -c1-300 > gpurun_out/repro_r2ab.txt; cat gpurun_out/repro_r2ab.txt
