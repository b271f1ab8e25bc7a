#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/two_stream_probe.py > gpurun_out/two_stream_r2c.txt 2> gpurun_out/two_stream_r2c.err; echo "exit=$?"
AQUA_PDL=0 timeout 600 python tools/two_stream_probe.py >> gpurun_out/two_stream_r2c.txt 2>> gpurun_out/two_stream_r2c.err; echo "exit=$?"
cat gpurun_out/two_stream_r2c.txt; tail -3 gpurun_out/two_stream_r2c.err
timeout 1200 python -m pytest tests/test_parity_b64.py -m gpu -q --tb=short > gpurun_out/parity_b64_r2c.log 2>&1; echo "parity exit=$? $(tail -1 gpurun_out/parity_b64_r2c.log)"
