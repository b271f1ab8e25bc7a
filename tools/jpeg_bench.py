"""JPEG tile ingest measured on the box: python tools/jpeg_bench.py [--tiles 64] [--size 1024]
  * host: cv2.imdecode per tile (what upstream LoadImages does) vs the library's Huffman-only parser, one thread and
    `--threads` threads (tiles/s);
  * device: aq_jpeg_reconstruct over a 64-tile batch with coefficients resident in HBM (CUDA events): idct + colour
    kernels against their algorithmic bytes (coefficients 2 B + planes 1 B written / read + 3 B per pixel out).
Prints one JSON line."""
import argparse
import ctypes as C
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import cv2
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aquaculture_b200 import engine, jpeg, tiles  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--tiles", type=int, default=64)
ap.add_argument("--size", type=int, default=1024)
ap.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()
cv2.setNumThreads(1)
rng = np.random.default_rng(0)


def scene(i):
    s = a.size
    base = rng.integers(0, 256, (s // 16 + 2, s // 16 + 2, 3), dtype=np.uint8)
    img = cv2.resize(base, (s, s), interpolation=cv2.INTER_CUBIC).astype(np.int32) + rng.integers(-12, 12, (s, s, 3))
    return np.clip(img, 0, 255).astype(np.uint8)


blobs = [cv2.imencode(".jpg", scene(i), [cv2.IMWRITE_JPEG_QUALITY, 75])[1] for i in range(a.tiles)]
info = jpeg.probe(blobs[0])
coef = torch.empty(a.tiles, info.coef_count, dtype=torch.int16).pin_memory()
quant = torch.empty(a.tiles, 192, dtype=torch.uint16).pin_memory()
cn, qn = coef.numpy(), quant.numpy()


def t_host(fn, threads):
    t0 = time.perf_counter()
    if threads == 1:
        for i in range(a.tiles):
            fn(i)
    else:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(fn, range(a.tiles)))
    return a.tiles / (time.perf_counter() - t0)


res = {"tiles": a.tiles, "size": a.size, "jpeg_bytes_per_tile": int(np.mean([b.size for b in blobs])), "threads": a.threads}
res["cv2_imdecode_1t"] = round(t_host(lambda i: cv2.imdecode(blobs[i], cv2.IMREAD_COLOR), 1), 1)
res["huffman_only_1t"] = round(t_host(lambda i: jpeg.entropy_decode(blobs[i], cn[i], qn[i]), 1), 1)
res["cv2_imdecode_mt"] = round(t_host(lambda i: cv2.imdecode(blobs[i], cv2.IMREAD_COLOR), a.threads), 1)
res["huffman_only_mt"] = round(t_host(lambda i: jpeg.entropy_decode(blobs[i], cn[i], qn[i]), a.threads), 1)

ctx = engine.Context(0)
dcoef, dquant = coef.cuda(), quant.cuda()
out = torch.empty(a.tiles, info.height, info.width, 3, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()


def run(src_dev):
    ctx.call("aq_jpeg_reconstruct", C.c_void_p(dcoef.data_ptr() if src_dev else coef.data_ptr()),
             C.c_void_p(dquant.data_ptr() if src_dev else quant.data_ptr()), a.tiles, info.width, info.height, info.ncomp, info.hs,
             1, int(src_dev), C.c_void_p(out.data_ptr()), None)


for src_dev in (1, 0):
    for _ in range(3):
        run(src_dev)
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(a.iters):
        run(src_dev)
    ctx.sync()
    ms = (time.perf_counter() - t0) * 1e3 / a.iters
    px = a.tiles * info.width * info.height
    samples = info.coef_count * a.tiles
    alg = samples * 2 + samples + samples + px * 3   # coefficients in, planes out, planes in, pixels out
    key = "device_resident" if src_dev else "device_from_pinned_host"
    res[key] = {"ms_per_batch": round(ms, 3), "tiles_per_s": round(a.tiles / ms * 1e3, 1), "algorithmic_gbs": round(alg / ms / 1e6, 1),
                "algorithmic_bytes": int(alg)}
ref = cv2.imdecode(blobs[0], cv2.IMREAD_COLOR)
res["bit_exact_vs_cv2"] = bool(np.array_equal(out[0].cpu().numpy(), ref))
ctx.close()
print(json.dumps(res))
