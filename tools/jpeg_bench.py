"""JPEG tile ingest measured on the box: python tools/jpeg_bench.py [--tiles 64] [--size 1024]
  * host: cv2.imdecode per tile (what upstream LoadImages does) vs the library's Huffman-only parser (packed output), one
    thread and `--threads` threads (tiles/s);
  * device: aq_jpeg_reconstruct over a 64-tile batch with dense coefficients resident in HBM, and
    aq_jpeg_reconstruct_packed from pinned host memory (copy stream, two batches in flight), CUDA events;
  * device, all of it: aq_jpeg_decode_device (parallel Huffman decoding + reconstruction) from the prepared scans in pinned
    host memory;
  * whole path (--e2e): .jpeg bytes -> detections through TileDetector, host Huffman on the io threads overlapped with the
    GPU, against the same detector fed cv2.imdecode pixels.
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
from aquaculture_b200 import engine, jpeg, weights  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--tiles", type=int, default=64)
ap.add_argument("--size", type=int, default=1024)
ap.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--e2e", type=int, default=0, help="batches of the whole-path comparison (0 = skip)")
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
N = a.tiles


def t_host(fn, threads, reps=2):
    best = 0.0
    for _ in range(reps):
        t0 = time.perf_counter()
        if threads == 1:
            for i in range(N):
                fn(i)
        else:
            with ThreadPoolExecutor(threads) as ex:
                list(ex.map(fn, range(N)))
        best = max(best, N / (time.perf_counter() - t0))
    return round(best, 1)


res = {"tiles": N, "size": a.size, "jpeg_bytes_per_tile": int(np.mean([b.size for b in blobs])), "threads": a.threads}
res["host_tiles_per_s"] = {
    "cv2_imdecode_1t": t_host(lambda i: cv2.imdecode(blobs[i], cv2.IMREAD_COLOR), 1),
    "huffman_packed_1t": t_host(lambda i: engine.entropy_decode_file(blobs[i]), 1),
    "cv2_imdecode_mt": t_host(lambda i: cv2.imdecode(blobs[i], cv2.IMREAD_COLOR), a.threads),
    "huffman_packed_mt": t_host(lambda i: engine.entropy_decode_file(blobs[i]), a.threads),
    "scan_prepare_1t": t_host(lambda i: engine.prepare_jpeg_file(blobs[i]), 1),
    "scan_prepare_mt": t_host(lambda i: engine.prepare_jpeg_file(blobs[i]), a.threads)}

ctx = engine.Context(0)
tl = [engine.entropy_decode_file(b) for b in blobs]
total = sum(t.entries.size for t in tl)
res["packed_bytes_per_tile"] = int((total * 4 + N * (info.coef_count // 64) * 4) / N)
res["dense_bytes_per_tile"] = int(info.coef_count * 2)
coef = np.zeros((N, info.coef_count), np.int16)
quant = np.zeros((N, 192), np.uint16)
for i, b in enumerate(blobs):
    jpeg.entropy_decode(b, coef[i], quant[i])
dcoef, dquant = torch.from_numpy(coef).cuda(), torch.from_numpy(quant.view(np.int16)).cuda()
out = torch.empty(N, info.height, info.width, 3, dtype=torch.uint8, device="cuda")
pin = [tuple(torch.from_numpy(x).pin_memory() for x in (np.empty((N, info.coef_count // 64), np.int32), np.empty(total, np.int32),
                                                         np.empty(N, np.int64), np.empty((N, 192), np.int16))) for _ in range(2)]
for d, e, b, q in pin:
    engine.pack_jpeg_tiles(tl, d.numpy().view(np.uint32), e.numpy().view(np.uint32), b.numpy().view(np.uint64), q.numpy().view(np.uint16))
torch.cuda.synchronize()


def run_dense():
    ctx.call("aq_jpeg_reconstruct", C.c_void_p(dcoef.data_ptr()), C.c_void_p(dquant.data_ptr()), N, info.width, info.height, info.ncomp,
             info.hs, 1, 1, C.c_void_p(out.data_ptr()), None)


slot = [0]


def run_packed():
    d, e, b, q = pin[slot[0]]
    slot[0] ^= 1
    ctx.call("aq_jpeg_reconstruct_packed", C.c_void_p(d.data_ptr()), C.c_void_p(e.data_ptr()), C.c_void_p(b.data_ptr()), total,
             C.c_void_p(q.data_ptr()), N, info.width, info.height, info.ncomp, info.hs, 1, C.c_void_p(out.data_ptr()), None)


sl = [engine.prepare_jpeg_file(b) for b in blobs]
scan_total = sum(t.scan.size for t in sl)
spin = []
for _ in range(2):
    sc = torch.empty(scan_total, dtype=torch.uint8).pin_memory()
    off, ln = torch.empty(N, dtype=torch.int64).pin_memory(), torch.empty(N, dtype=torch.int32).pin_memory()
    tab, qq = torch.empty(N, 8640, dtype=torch.uint8).pin_memory(), torch.empty(N, 192, dtype=torch.int16).pin_memory()
    stt = torch.zeros(N, dtype=torch.int32).pin_memory()
    engine.pack_jpeg_scans(sl, sc.numpy(), off.numpy().view(np.uint64), ln.numpy().view(np.uint32), tab.numpy(), qq.numpy().view(np.uint16))
    spin.append((sc, off, ln, tab, qq, stt))
sslot = [0]


def run_all_device():
    sc, off, ln, tab, qq, stt = spin[sslot[0]]
    sslot[0] ^= 1
    ctx.call("aq_jpeg_decode_device", C.c_void_p(sc.data_ptr()), C.c_void_p(off.data_ptr()), C.c_void_p(ln.data_ptr()), scan_total,
             C.c_void_p(tab.data_ptr()), C.c_void_p(qq.data_ptr()), N, info.width, info.height, info.ncomp, info.hs, 1,
             C.c_void_p(out.data_ptr()), C.c_void_p(stt.data_ptr()), None)


px = N * info.width * info.height
samples = info.coef_count * N
for key, fn, in_bytes in (("device_dense_resident", run_dense, samples * 2), ("device_packed_from_pinned_host", run_packed, res["packed_bytes_per_tile"] * N),
                          ("all_device_from_pinned_host", run_all_device, scan_total + samples * 2 * 2)):
    for _ in range(3):
        fn()
    ctx.sync()
    t0 = time.perf_counter()
    for _ in range(a.iters):
        fn()
    ctx.sync()
    ms = (time.perf_counter() - t0) * 1e3 / a.iters
    alg = in_bytes + samples + samples + px * 3   # coefficients in, planes out, planes in, pixels out
    res[key] = {"ms_per_batch": round(ms, 3), "tiles_per_s": round(N / ms * 1e3, 1), "algorithmic_gbs": round(alg / ms / 1e6, 1),
                "algorithmic_bytes": int(alg)}
    res[key]["bit_exact_vs_cv2"] = bool(np.array_equal(out[0].cpu().numpy(), cv2.imdecode(blobs[0], cv2.IMREAD_COLOR)))
ctx.close()

if a.e2e > 0:
    det = engine.TileDetector(weights.synthetic_state(3, 0), 3, batch=N, imgsz=(640, 640), device=0)
    pool = ThreadPoolExecutor(a.threads)

    def via_jpeg():
        tiles_ = list(pool.map(engine.entropy_decode_file, blobs))
        d, e, b, q = det.jpeg_staging(N, info.coef_count, sum(t.entries.size for t in tiles_))
        tot = engine.pack_jpeg_tiles(tiles_, d, e, b, q, pool=pool)
        return det.detect_submit_jpeg(d, e, b, tot, q, info.width, info.height, info.ncomp, info.hs)

    def via_scans():
        (sc, off, ln, tab, q, stt), tot, i0 = engine.prepare_jpeg_batch(det, blobs, n_threads=a.threads)
        return det.detect_submit_jpeg_scans(sc, off, ln, tot, tab, q, stt, i0.width, i0.height, i0.ncomp, i0.hs)

    def via_cv2():
        batch = det.pinned_input(N, info.height, info.width)
        bn = batch.numpy()

        def one(i):
            bn[i] = cv2.imdecode(blobs[i], cv2.IMREAD_COLOR)
        list(pool.map(one, range(N)))
        return det.detect_submit(batch, bgr=True)

    counts = {}
    for name, submit in (("jpeg_all_device", via_scans), ("jpeg_device_decode", via_jpeg), ("cv2_host_decode", via_cv2)):
        det.detect_wait(submit())
        t0 = time.perf_counter()
        prev = None
        for _ in range(a.e2e):
            tk = submit()
            if prev is not None:
                det.detect_wait(prev)
            prev = tk
        r = det.detect_wait(prev)
        counts[name] = r.counts.copy()
        res.setdefault("whole_path_tiles_per_s", {})[name] = round(a.e2e * N / (time.perf_counter() - t0), 1)
    res["whole_path_tiles_per_s"]["identical_detection_counts"] = bool(np.array_equal(counts["jpeg_device_decode"], counts["cv2_host_decode"]) and
                                                                       np.array_equal(counts["jpeg_all_device"], counts["cv2_host_decode"]))
    det.close()
print(json.dumps(res))
