#!/bin/bash
# the allocation-free entry point: its test alone in a fresh process, then the whole GPU suite (which ends with it)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone exit=$? $(tail -1 gpurun_out/pytest_noalloc_alone.log)"; grep -E "Error|FAILED|assert" gpurun_out/pytest_noalloc_alone.log | cut -c1-300 | head
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_noalloc_all.log 2>&1; echo "all exit=$? $(tail -1 gpurun_out/pytest_noalloc_all.log)"; grep -E "Error|FAILED" gpurun_out/pytest_noalloc_all.log | cut -c1-300 | head
