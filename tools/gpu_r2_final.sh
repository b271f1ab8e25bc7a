#!/bin/bash
# final round-2 check: whole GPU suite, smoke(), reproducibility probe, the default bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -q -m gpu > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest -m gpu exit=$? $(tail -1 gpurun_out/pytest_gpu_final.log)"; grep -E "Error|FAILED" gpurun_out/pytest_gpu_final.log | cut -c1-400 | head
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke exit=$? $(tail -1 gpurun_out/smoke_final.log | cut -c1-200)"
timeout 600 python tools/repro_probe.py --iters 1000 2>> gpurun_out/repro_final.err | grep -v "^[0-9]* conv\|^[0-9]* detect\|^[0-9]* sppf" | cut -c1-300 > gpurun_out/repro_final.txt; cat gpurun_out/repro_final.txt
timeout 900 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench exit=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_final.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline'], d['parity']['ok'], d['parity'].get('within_tolerance'), d['clocks'], d['cpu_baseline'])"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_final_ref.json 2> gpurun_out/bench_final_ref.err; echo "ref exit=$?"; cut -c1-600 gpurun_out/bench_final_ref.json
