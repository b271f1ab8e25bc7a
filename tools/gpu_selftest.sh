#!/bin/bash
# Runs the GPU parity suite in separate processes (a trapped kernel poisons its CUDA context) and keeps logs.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name, pytest -k expr
  timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_jpeg.py -m gpu -q -x --tb=short -k "$2" > gpurun_out/selftest_$1.log 2>&1
  echo "$1 exit=$? $(tail -1 gpurun_out/selftest_$1.log)"
}
run conv1x1 "conv1x1 and not folded"
run folded "folded_upsample"
run conv3x3s1 "conv3x3_s1 and not variants and not dense or residual"
run dense "dense_taps"
run variants "s1_variants"
run s2variants "s2_variants"
run mblk "mblk_variants or fast_silu"
run conv3x3s2 "conv3x3_s2"
run cluster "cluster"
run persistent "many_tiles"
run detect "detect_decode"
run prepost "preprocess or sppf or tile_stats"
run nms "nms"
run e2e "e2e or label_files"
run jpeg "jpeg"
run rest "not (conv1x1 or folded_upsample or conv3x3_s1 or residual or dense_taps or s1_variants or s2_variants or mblk_variants or fast_silu or conv3x3_s2 or cluster or many_tiles or detect_decode or preprocess or sppf or tile_stats or nms or e2e or label_files or jpeg)"
