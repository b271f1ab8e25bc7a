"""Sweep of the depth-first stage configuration (engine.stage_plan): conv-stack time per 64-tile batch for each
AQUA_STAGES spec, same weights and tiles, predictions compared with the unstaged plan.
  python tools/stage_sweep.py [--batch 64] [--img 640] [--iters 30] spec [spec ...]     (spec "0" = unstaged)"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from aquaculture_b200 import engine, tiles, weights

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--img", type=int, default=640)
ap.add_argument("--iters", type=int, default=30)
ap.add_argument("--rounds", type=int, default=2)
ap.add_argument("specs", nargs="+")
a = ap.parse_args()
sd = weights.synthetic_state(3, 0)
t = torch.from_numpy(tiles.synthetic_tiles(range(a.batch), a.img, a.img)).pin_memory()
dets, ref = {}, None
for spec in a.specs:
    os.environ["AQUA_STAGES"] = spec
    t0 = time.perf_counter()
    det = engine.TileDetector(sd, 3, batch=a.batch, imgsz=(a.img, a.img))
    det.detect(t)
    pred = det.ctx.pred()
    if ref is None:
        ref = pred
    dets[spec] = (det, float(np.abs(pred - ref)[..., 4:].max()), time.perf_counter() - t0)
best = {s: 1e9 for s in a.specs}
for r in range(a.rounds):
    for spec in a.specs:
        det = dets[spec][0]
        for _ in range(5):
            det.ctx.forward()
        det.ctx.sync()
        t0 = time.perf_counter()
        for _ in range(a.iters):
            det.ctx.forward()
        det.ctx.sync()
        best[spec] = min(best[spec], (time.perf_counter() - t0) / a.iters * 1e3)
for spec in a.specs:
    det, err, plan_s = dets[spec]
    print(f"stages {spec:28s} {best[spec]:7.3f} ms/batch  {a.batch / best[spec] * 1e3:8.0f} tiles/s  launches {det.n_launches:4d}  "
          f"max|conf - first spec| {err:.2e}  plan {plan_s:.0f}s  {det.stages}", flush=True)
for d in dets.values():
    d[0].close()
