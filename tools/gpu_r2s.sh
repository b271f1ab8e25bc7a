#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref exit=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_r2_n$N.json 2> gpurun_out/bench_r2_n$N.err; echo "bench n$N exit=$?"

python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r2_n$N.json').read())
for k in ('value','n_gpus','ms_per_step','e2e','parity','jpeg_ingest','counters','clocks'):
    v=d.get(k)
    if isinstance(v,dict): v={a:b for a,b in v.items() if not isinstance(b,(str,dict)) or (isinstance(b,str) and len(b)<50)}
    print(k, v)
PY
