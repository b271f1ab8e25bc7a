#!/bin/bash
# CLI over two GPUs (spawned ranks) against one GPU on the same synthetic tiles: identical label files
mkdir -p gpurun_out
rm -rf /tmp/runs
timeout 600 python detect.py --weights synthetic --synthetic 768 --nosave --save-txt --save-conf --project /tmp/runs --name one --device 0 2>&1 | tail -4
timeout 600 python detect.py --weights synthetic --synthetic 768 --nosave --save-txt --save-conf --project /tmp/runs --name two --device 0,1 2>&1 | tail -4
echo "exit codes ok; files: $(ls /tmp/runs/one/labels | wc -l) $(ls /tmp/runs/two/labels | wc -l)"
diff -rq /tmp/runs/one/labels /tmp/runs/two/labels && echo "label files identical"
