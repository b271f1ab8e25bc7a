#!/bin/bash
# the per-launch ncu pass of tools/gpu_r2_ncu.sh alone (DRAM traffic file for bench.py's roofline.traffic): needed whenever the
# kernel source id changes without the conv kernels themselves changing
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.sum,sm__ops_path_tensor_op_utchmma_src_fp16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tensor.sum,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,lts__t_bytes.sum
timeout 300 python tools/forward_only.py > gpurun_out/fwd_plain.log 2>&1 &&
timeout 1500 ncu --metrics $M --clock-control none -k regex:'conv_igemm|sppf' -s 216 -c 72 --csv --log-file gpurun_out/layers_r2.csv \
    python tools/forward_only.py > gpurun_out/ncu_layers.log 2>&1
echo "ncu layers exit=$?"; tail -2 gpurun_out/fwd_plain.log
timeout 300 python tools/profile_ops.py --out gpurun_out/ops_r2.txt > /dev/null 2> gpurun_out/ops_r2.err; tail -1 gpurun_out/ops_r2.txt
