#!/bin/bash
# round 2 ncu evidence (one gpurun call): per-launch table of one forward pass of the benchmarked plan (duration, DRAM bytes,
# tensor-path metrics that are meaningful for UTCHMMA) + --set full captures of four representative convs.
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.sum,sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tensor.sum,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_bytes.sum
timeout 300 python tools/forward_only.py > gpurun_out/fwd_plain.log 2>&1 &&
timeout 1500 ncu --metrics $M --clock-control none -k regex:'conv_igemm|sppf' -s 216 -c 72 --csv --log-file gpurun_out/layers_r2.csv \
    python tools/forward_only.py > gpurun_out/ncu_layers.log 2>&1
echo "ncu layers exit=$?"; cat gpurun_out/fwd_plain.log | tail -2
timeout 300 python tools/profile_ops.py --out gpurun_out/ops_r2.txt > /dev/null 2> gpurun_out/ops_r2.err; tail -1 gpurun_out/ops_r2.txt
for spec in "stem 142" "l1_chain_cv12 143" "c192_3x3_40 162" "pair_c1536_1x1_20 182" "c384_3x3_20 210"; do
  set -- $spec
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_igemm -s $2 -c 1 -f -o gpurun_out/full_r2_$1 \
      python tools/forward_only.py > gpurun_out/ncu_full_$1.log 2>&1
  echo "ncu full $1 exit=$?"
done
ls -la gpurun_out/*.ncu-rep
