#!/bin/bash
# round-2 wrap-up after a kernel change: whole GPU suite, smoke, per-launch ncu pass (DRAM traffic file), op table
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_wrap.log 2>&1; echo "pytest -m gpu exit=$? $(tail -1 gpurun_out/pytest_gpu_wrap.log)"; grep -E "Error|FAILED" gpurun_out/pytest_gpu_wrap.log | cut -c1-300 | head
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_wrap.log 2>&1; echo "smoke exit=$? $(tail -1 gpurun_out/smoke_wrap.log | cut -c1-160)"
bash tools/gpu_r2_ncu_layers.sh
