"""Smallest program that launches every kernel family once, for compute-sanitizer (memcheck / racecheck): a batch-2 640x640
detector (FLAT / HALO / HALO2 / TAP feeds, resident and streamed weights, Detect epilogue + candidate mask, SPPF pool,
preprocess, NMS + mask kernel), a thread-block-cluster conv, 1024^2 -> 640 letterbox, raster windows + tile stats, and the
all-device JPEG decoder.   compute-sanitizer --tool memcheck python tools/sanitize_prog.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2, numpy as np, torch
from aquaculture_b200 import engine, tiles, weights
sd = weights.synthetic_state(3, 0)
det = engine.TileDetector(sd, 3, batch=2, imgsz=(640, 640))
r = det.detect(torch.from_numpy(tiles.synthetic_tiles(range(2), 640, 640)).pin_memory())
print("detect 640:", r.counts.tolist())
r = det.detect(torch.from_numpy(tiles.synthetic_tiles(range(2), 1024, 1024)).pin_memory())
print("detect 1024->640:", r.counts.tolist())
pred = torch.from_numpy(det.ctx.pred()).cuda()
d, c, k = engine.nms_device(det.ctx, pred, classes=[0, 2])   # external predictions: nms_mask_kernel
print("nms (mask kernel):", c.tolist())
raster = torch.from_numpy(tiles.synthetic_tiles([7], 1280, 1920)[0]).cuda()
r = det.detect_windows(raster, 640, 0, 2)
print("windows:", r.counts.tolist(), engine.tile_stats(det.ctx, raster, tile=640, win0=0, n=2).tolist())
blobs = [cv2.imencode(".jpg", tiles.synthetic_tiles([i], 640, 640)[0], [cv2.IMWRITE_JPEG_QUALITY, 75])[1] for i in range(2)]
(sc, off, ln, tab, q, stt), tot, i0 = engine.prepare_jpeg_batch(det, blobs, n_threads=2)
r = det.detect_wait(det.detect_submit_jpeg_scans(sc, off, ln, tot, tab, q, stt, i0.width, i0.height, i0.ncomp, i0.hs))
print("jpeg:", r.counts.tolist(), stt.tolist())
det.close()
# a thread-block-cluster conv (weight tiles shared by TMA multicast) and a dense-tap stem variant
ctx = engine.Context(0)
ctx.plan_begin(2, 64, 64, 3)
a, b = ctx.buffer(40, 40, 192), ctx.buffer(40, 40, 192)
g = torch.Generator().manual_seed(0)
ctx.conv((a, 0), (b, 0), torch.randn(192, 192, 3, 3, generator=g) * 0.02, torch.zeros(192), 3, 1, 2, cluster=2)
ctx.plan_end()
ctx.forward(); ctx.sync()
print("cluster conv ok")
ctx.close()
