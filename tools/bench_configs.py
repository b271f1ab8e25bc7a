#!/usr/bin/env python
"""The BASELINE.json configurations bench.py does not time (bench.py = configs[1]); one JSON line each.

  python tools/bench_configs.py [--which 3,4,5,raster,stats] [--tiles 200000]
  python -m torch.distributed.run --nproc-per-node N ... tools/bench_configs.py --which 3      (sharded sweep)

  3      BD-Ortho-scale sweep: --tiles synthetic 640^2 tiles strided over the ranks, device-resident tiles
         (4 rotating batches), preprocess + conv stack + decode + NMS per batch of 64; tiles/s, max over ranks
  4      NMS stress: aq_nms alone on (64, 25200, 8) fp32 with ~20 000 rows above conf per image, max_det 1000;
         images/s, DRAM GB/s of the algorithmic bytes, candidate pairs/s
  5      --img 1280, batch 16 fp16: tiles/s and conv-stack TFLOP/s (191.5 GFLOP per tile)
  raster tiling fused into inference: 36 windows of 1024^2 of a device-resident 6144^2 raster -> 640^2
  stats  blank-tile statistics kernel on 64 tiles of 1024^2: GB/s vs the HBM peak
All timings: CUDA events on the launching stream after warm-up; inputs larger than L2 or rotated.
"""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from aquaculture_b200 import arch, engine, tiles, weights

NC = 3


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "measured"
    except Exception:
        return 6650.0, 1400.0, "fallback"


def timed(stream, fn, reps, warm=3):
    with torch.cuda.stream(stream):
        for _ in range(warm):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            fn()
        e1.record()
    stream.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--which", default="3,4,5,raster,stats")
    ap.add_argument("--tiles", type=int, default=200000)
    a = ap.parse_args()
    which = a.which.split(",")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    hbm, tf_peak, src = peaks()
    sd = weights.synthetic_state(NC, 0)
    stream = torch.cuda.Stream()
    sarg = C.c_void_p(stream.cuda_stream)

    def emit(d):
        if rank == 0:
            print(json.dumps(d), flush=True)

    if "3" in which:
        B = 64
        det = engine.TileDetector(sd, NC, batch=B, imgsz=(640, 640), device=local)
        sets = [torch.from_numpy(tiles.synthetic_tiles([rank + world * (s * B + i) for i in range(B)], 640, 640)).cuda() for s in range(4)]
        steps = (a.tiles // world + B - 1) // B
        for i in range(3):
            det.detect_wait(det.detect_submit(sets[i % 4], bgr=True, src_is_device=True, stream=sarg))
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_det = 0
        with torch.cuda.stream(stream):
            e0.record()
            last = None
            for i in range(steps):
                t = det.detect_submit(sets[i % 4], bgr=True, src_is_device=True, stream=sarg)
                if last is not None:
                    n_det += int(det.detect_wait(last).counts.sum())
                last = t
            n_det += int(det.detect_wait(last).counts.sum())
            e1.record()
        stream.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            c = torch.tensor([n_det], device="cuda", dtype=torch.int64)
            dist.all_reduce(c)
            n_det = int(c.item())
        emit({"config": "configs[2]: sweep of synthetic 640x640 tiles strided over the ranks, device-resident tiles",
              "tiles": steps * B * world, "n_gpus": world, "seconds": round(ms / 1e3, 3),
              "value": round(steps * B * world / (ms / 1e3), 1), "unit": "tiles/s", "detections": n_det, "batch_per_gpu": B})
        det.close()

    if "4" in which and rank == 0:
        ctx = engine.Context(local)
        B, rows, npass = 64, 25200, 20000
        g = torch.Generator().manual_seed(11)
        pred = torch.zeros(B, rows, 5 + NC)
        pred[..., 0:2] = torch.rand(B, rows, 2, generator=g) * 640
        pred[..., 2:4] = torch.rand(B, rows, 2, generator=g) * 40 + 4
        obj = torch.rand(B, rows, generator=g) * 0.2
        idx = torch.rand(B, rows, generator=g).argsort(1)[:, :npass]
        obj.scatter_(1, idx, torch.rand(B, npass, generator=g) * 0.7 + 0.3)
        pred[..., 4] = obj
        pred[..., 5:] = torch.rand(B, rows, NC, generator=g)
        pd = pred.cuda().contiguous()
        dets = torch.zeros(B, 1000, 6, device="cuda")
        cnt = torch.zeros(B, dtype=torch.int32, device="cuda")
        keep = torch.zeros(B, 1000, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        args = (C.c_void_p(pd.data_ptr()), B, rows, 5 + NC, C.c_float(0.25), C.c_double(0.45), 1000, 0, None, 0, C.c_float(7680.0),
                30000, C.c_void_p(dets.data_ptr()), C.c_void_p(cnt.data_ptr()), C.c_void_p(keep.data_ptr()), sarg)
        ms = timed(stream, lambda: ctx.call("aq_nms", *args), 10)
        conf = pred[..., 4:5] * pred[..., 5:]
        ncand = int(((pred[..., 4] > 0.25) & (conf.max(-1).values > 0.25)).sum())
        per_img = ncand / B
        bytes_alg = B * rows * (5 + NC) * 4 + int(cnt.sum()) * 24
        emit({"config": "configs[3]: NMS stress, (64, 25200, 8) fp32, ~20k rows above conf per image, 3 classes, max_det 1000",
              "ms_per_batch": round(ms, 3), "value": round(B / (ms / 1e3), 1), "unit": "images/s",
              "candidates_per_image": round(per_img, 1), "kept_per_image": round(float(cnt.float().mean()), 1),
              "algorithmic_gbs": round(bytes_alg / ms / 1e6, 1), "frac_of_hbm_peak": round(bytes_alg / ms / 1e6 / hbm, 4),
              "candidate_pairs_per_s": round(B * per_img * per_img / 2 / 3 / (ms / 1e3), 0),
              "note": "pairs = n^2/2 within a class (3 classes); keep indices bit-exact vs torchvision in tests/test_gpu_kernels.py::test_nms_stress_20k"})
        ctx.close()

    if "5" in which and rank == 0:
        B, S = 16, 1280
        det = engine.TileDetector(sd, NC, batch=B, imgsz=(S, S), device=local)
        sets = [torch.from_numpy(tiles.synthetic_tiles(range(s * B, s * B + B), S, S)).cuda() for s in range(4)]
        k = [0]

        def step():
            det.detect_wait(det.detect_submit(sets[k[0] % 4], bgr=True, src_is_device=True, stream=sarg))
            k[0] += 1
        ms = timed(stream, step, 20)
        ms_fwd = timed(stream, lambda: det.ctx.call("aq_forward", sarg), 20)
        flop = arch.conv_flops_per_tile(S, S, NC) * B
        emit({"config": "configs[4]: --img 1280, batch 16 fp16, synthetic 1280x1280 tiles, conf 0.25 / IoU 0.45",
              "ms_per_batch": round(ms, 3), "value": round(B / (ms / 1e3), 1), "unit": "tiles/s",
              "equiv_640_tiles_per_s": round(4 * B / (ms / 1e3), 1), "ms_conv_stack": round(ms_fwd, 3),
              "conv_tflops": round(flop / ms_fwd / 1e9, 1), "frac_of_sustained_peak": round(flop / ms_fwd / 1e9 / tf_peak, 4),
              "peak_source": src, "detect_rows": 100800})
        det.close()

    if "raster" in which and rank == 0:
        T, G = 1024, 6
        det = engine.TileDetector(sd, NC, batch=36, imgsz=(640, 640), device=local)
        raster = torch.randint(0, 256, (T * G, T * G, 3), dtype=torch.uint8, device="cuda")
        det.detect_windows(raster, T, 0, 36)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        e0.record()
        for _ in range(reps):
            det.detect_windows(raster, T, 0, 36)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        emit({"config": "8f-3: 36 windows of 1024^2 cut from a device-resident 6144^2 raster, letterboxed to 640^2, detect + NMS + D2H",
              "ms_per_raster": round(ms, 3), "value": round(36 / (ms / 1e3), 1), "unit": "tiles/s", "batch": 36})
        det.close()

    if "stats" in which and rank == 0:
        ctx = engine.Context(local)
        t = torch.randint(0, 256, (64, 1024, 1024, 3), dtype=torch.uint8, device="cuda")
        out = np.zeros((64, 4), np.int32)
        torch.cuda.synchronize()
        fn = lambda: ctx.call("aq_tile_stats", C.c_void_p(t.data_ptr()), 64, 1024, 1024, 1, 0, 0, 0, out.ctypes.data, sarg)
        ms = timed(stream, fn, 10)
        ms_k = timed(stream, lambda: ctx.call("aq_tile_stats", C.c_void_p(t.data_ptr()), 64, 1024, 1024, 1, 0, 0, 0, None, sarg), 20)
        b = 64 * 1024 * 1024 * 3
        emit({"config": "8f-5: blank-tile statistics of 64 tiles of 1024^2 (201 MB read); kernels alone (memset + init + stats + finish) "
                        "and the whole call incl. D2H of 64 x 16 B and the stream sync",
              "ms_kernels": round(ms_k, 4), "achieved_gbs": round(b / ms_k / 1e6, 1), "frac_of_hbm_peak": round(b / ms_k / 1e6 / hbm, 3),
              "ms_per_call": round(ms, 3)})
        ctx.close()

    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
