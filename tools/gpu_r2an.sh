#!/bin/bash
# attribution of the high-resolution layers (current build): AQUA_DBG 1 = no TMA stores, 2 = no epilogue math / staging, 8 = no TMEM loads,
# 32 = no staging handshake.  Tilings pinned to the shipped choice of each layer so that only the switched-off part changes.
mkdir -p gpurun_out
export AQUA_TUNE_LOG=0
{
for spec in "64 320 320 16 48 3 1 --mblk 4 --mode 1" "64 160 160 48 48 1 1" "64 160 160 48 48 3 1 --res" "64 160 160 96 96 1 1" "64 80 80 96 96 3 1 --res" "64 80 80 96 96 1 1"; do
  for dbg in 0 1 2 3 8 11; do   # never 32: it deadlocks the two-set epilogue until the bounded waits trap (minutes per run)
    echo -n "dbg=$dbg  "; AQUA_DBG=$dbg timeout 120 python tools/one_op.py $spec 2>&1 | tail -1 | cut -c1-200
  done
done
} > gpurun_out/attribution_r2an.txt
cat gpurun_out/attribution_r2an.txt | cut -c1-150
