"""NMS on BASELINE.json configs[3] (~20 k candidates per image): time of aq_nms per 64 images (CUDA events) and, with
AQUA_NMS_CLK=1 in the environment, the phase timeline of image 0's CTA.  python tools/nms_stress.py [--iters 10]"""
import argparse, ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from aquaculture_b200 import engine, tiles, weights
ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--shift", type=float, default=10.0)
a = ap.parse_args()
B, IMG, NC = 64, 640, 3
det = engine.TileDetector(weights.synthetic_state(NC, 0, obj_shift=a.shift), NC, batch=B, imgsz=(IMG, IMG))
t = torch.from_numpy(tiles.synthetic_tiles(range(B), IMG, IMG, seed=0)).cuda()
for _ in range(2):
    det.detect(t, bgr=True, src_is_device=True)
pred = torch.from_numpy(det.ctx.pred())
cand = ((pred[..., 4] > 0.25) & ((pred[..., 5:] * pred[..., 4:5]).amax(-1) > 0.25)).sum(1)
rows = pred.shape[1]
dets_d = torch.zeros(B, det.max_det, 6, device="cuda")
cnt_d = torch.zeros(B, dtype=torch.int32, device="cuda")
args = (None, B, rows, 5 + NC, C.c_float(0.25), C.c_double(0.45), det.max_det, 0, None, 0, C.c_float(7680.0), 30000,
        C.c_void_p(dets_d.data_ptr()), C.c_void_p(cnt_d.data_ptr()), None, None)
det.ctx.call("aq_nms", *args)
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(a.iters):
    det.ctx.call("aq_nms", *args)
torch.cuda.synchronize()
det.ctx.sync()
ms = (time.perf_counter() - t0) / a.iters * 1e3
print(f"candidates per image: mean {float(cand.float().mean()):.0f} min {int(cand.min())} max {int(cand.max())}; kept per image {float(cnt_d.float().mean()):.1f}; "
      f"aq_nms {ms:.4f} ms per {B} images (wall clock around {a.iters} calls + sync)")
det.close()
