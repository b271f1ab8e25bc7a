#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q --tb=short -k "nms or e2e or label or cli" > gpurun_out/nms_r2t.log 2>&1; echo "nms/e2e tests exit=$? $(tail -1 gpurun_out/nms_r2t.log)"; grep -E "Error|FAILED" gpurun_out/nms_r2t.log | cut -c1-300 | head
timeout 900 python -m pytest tests/test_parity_b64.py tests/test_jpeg.py -m gpu -q --tb=short > gpurun_out/par_r2t.log 2>&1; echo "parity/jpeg tests exit=$? $(tail -1 gpurun_out/par_r2t.log)"
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-jpeg --parity-tiles 64 > gpurun_out/bench_r2t.json 2> gpurun_out/bench_r2t.err; python -c "
import json; d=json.load(open('gpurun_out/bench_r2t.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['ms_per_conv_stack'], d['aux_kernels']['nms_kernel'], d['extra_configs']['nms_stress'], d['parity']['ok'], d['parity']['unmatched'])"
