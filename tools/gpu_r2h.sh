#!/bin/bash
mkdir -p gpurun_out
echo "baseline/_ref on the GPU box:"; ls -la baseline/_ref 2>&1 | head -5; find / -name "*.pt" -size +10M 2>/dev/null | grep -v -E "site-packages|/proc/" | head -5; find / -type d -name "yolov5*" 2>/dev/null | head -5; echo "(end of search)"
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_gpu_r2h.log 2>&1; echo "pytest -m gpu exit=$? $(tail -1 gpurun_out/pytest_gpu_r2h.log)"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r2h.log 2>&1; echo "smoke exit=$? $(tail -1 gpurun_out/smoke_r2h.log)"
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_r2h.json 2> gpurun_out/bench_r2h.err; echo "bench exit=$?"; cat gpurun_out/bench_r2h.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k in ('value','ms_per_step','e2e','roofline','parity','tilings','extra_configs','aux_kernels','jpeg_ingest','clocks'):
    v=d.get(k)
    if isinstance(v,dict): v={a:b for a,b in v.items() if not isinstance(b,str) or len(b)<60}
    print(k, v)
"
tail -3 gpurun_out/bench_r2h.err
