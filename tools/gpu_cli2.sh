#!/bin/bash
# CLI over two GPUs (spawned ranks) against one GPU on the same synthetic tiles: identical label files; hashes for cross-box comparison
mkdir -p gpurun_out
rm -rf /tmp/runs
python tools/weights_hash.py
timeout 600 python detect.py --weights synthetic --synthetic 768 --nosave --save-txt --save-conf --project /tmp/runs --name one --device 0 2>&1 | grep "labels saved"
timeout 600 python detect.py --weights synthetic --synthetic 768 --nosave --save-txt --save-conf --project /tmp/runs --name two --device 0,1 2>&1 | grep "labels saved"
for r in one two; do echo "$r $(cat /tmp/runs/$r/labels/*.txt | sha256sum | cut -c1-16) $(cat /tmp/runs/$r/labels/*.txt | wc -l)"; done
diff -rq /tmp/runs/one/labels /tmp/runs/two/labels && echo "label files identical"
lscpu | grep "Model name"
