"""A/B of plan-time environment switches on the benchmarked conv stack: for each "NAME=VALUE[,NAME=VALUE]" spec (or "-" for the
defaults) a batch-64 detector is planned under that environment; all are then timed interleaved (forward passes, wall clock
around N graph launches + sync) and their predictions compared with the first spec's.
  python tools/ab_env.py - AQUA_SERPENTINE=0 ..."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from aquaculture_b200 import engine, tiles, weights
specs = sys.argv[1:] or ["-"]
B, IMG = int(os.environ.get("AB_BATCH", 64)), int(os.environ.get("AB_IMG", 640))
sd = weights.synthetic_state(3, 0)
t = torch.from_numpy(tiles.synthetic_tiles(range(B), IMG, IMG)).pin_memory()
dets, ref = {}, None
for spec in specs:
    kv = dict(x.split("=", 1) for x in spec.split(",")) if spec != "-" else {}
    old = {k: os.environ.get(k) for k in kv}
    os.environ.update(kv)
    det = engine.TileDetector(sd, 3, batch=B, imgsz=(IMG, IMG))
    for k, v in old.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v
    det.detect(t)
    pred = det.ctx.pred()
    ref = pred if ref is None else ref
    dets[spec] = (det, float(np.abs(pred - ref).max()))
best = {s: 1e9 for s in specs}
for r in range(3):
    for spec in specs:
        det = dets[spec][0]
        for _ in range(5):
            det.ctx.forward()
        det.ctx.sync()
        t0 = time.perf_counter()
        for _ in range(40):
            det.ctx.forward()
        det.ctx.sync()
        best[spec] = min(best[spec], (time.perf_counter() - t0) / 40 * 1e3)
for spec in specs:
    print(f"{spec:40s} {best[spec]:7.3f} ms per conv stack  {B / best[spec] * 1e3:8.0f} tiles/s  max |pred - first| {dets[spec][1]:.2e}", flush=True)
for d in dets.values():
    d[0].close()
