#!/bin/bash
# 2-GPU check: bench under torchrun (both arms), CLI sharded over two devices
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref exit=$?"; cat gpurun_out/bench_ref_n2.json | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/bench_r2_n2.json 2> gpurun_out/bench_r2_n2.err; echo "bench n2 exit=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r2_n2.json').read())
for k in ('value','n_gpus','ms_per_step','e2e','parity','jpeg_ingest','extra_configs','counters'):
    v=d.get(k)
    if isinstance(v,dict): v={a:b for a,b in v.items() if not isinstance(b,str) or len(b)<50}
    print(k, v)
PY
tail -3 gpurun_out/bench_r2_n2.err
rm -rf /tmp/cli2; timeout 600 python detect.py --weights synthetic --synthetic 1024 --nosave --save-txt --save-conf --device 0,1 --project /tmp/cli2 --name run > gpurun_out/cli_n2.log 2>&1; echo "cli 2 devices exit=$?"; tail -4 gpurun_out/cli_n2.log; ls /tmp/cli2/run/labels | wc -l
rm -rf /tmp/cli1; timeout 600 python detect.py --weights synthetic --synthetic 1024 --nosave --save-txt --save-conf --device 0 --project /tmp/cli1 --name run > gpurun_out/cli_n1.log 2>&1; echo "cli 1 device exit=$?"; tail -4 gpurun_out/cli_n1.log
diff -rq /tmp/cli1/run/labels /tmp/cli2/run/labels && echo "label files of the 1-device and 2-device runs are identical"
