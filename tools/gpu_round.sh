#!/bin/bash
# One GPU-box visit: parity suite, bench, per-op table, ncu launch list, ncu --set full of one dominant conv.
# usage: tools/gpu_round.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out
bash tools/gpu_selftest.sh 2>&1 | tee gpurun_out/selftest_$TAG.txt
timeout 600 python bench.py --steps 30 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench exit=$?"; cat gpurun_out/bench_$TAG.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2>> gpurun_out/bench_$TAG.err; cat gpurun_out/bench_ref_$TAG.json
timeout 300 python tools/profile_ops.py --out gpurun_out/ops_$TAG.txt > /dev/null 2> gpurun_out/ops_$TAG.err; tail -1 gpurun_out/ops_$TAG.txt
if [ -z "$NO_NCU" ]; then
export AQUA_TUNE_CACHE=gpurun_out/tune_$TAG.txt   # the ncu pass replays the plain run's tilings: no tuning launches in the list
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_$TAG.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch_$TAG.log 2>&1
echo "ncu launches exit=$?"
unset AQUA_TUNE_CACHE
for shp in "64 80 80 96 96 3 1"; do
  n=$(echo $shp | tr ' ' '_')
  timeout 120 python tools/one_op.py $shp > gpurun_out/oneop_${n}_$TAG.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s 3 -c 1 -f -o gpurun_out/prof_${n}_$TAG \
      python tools/one_op.py $shp > gpurun_out/ncu_full_${n}_$TAG.log 2>&1
  echo "ncu full $n exit=$?"
done
fi
