"""A/B of plan-time environment switches on the benchmarked conv stack: for each "NAME=VALUE[,NAME=VALUE]" spec (or "-" for the
defaults; several assignments joined with "+") a batch-64 detector is planned under that environment; all are then timed interleaved (forward passes, wall clock
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
    kv = dict(x.split("=", 1) for x in spec.split("+")) if spec != "-" else {}  # NAME=VALUE+NAME=VALUE
    old = {k: os.environ.get(k) for k in kv}
    os.environ.update(kv)
    from aquaculture_b200 import _lib
    shipped = _lib.TUNE_TABLE
    if "TABLE" in kv:  # TABLE=<path>: another tiling table for this spec only (read when the context is created)
        _lib.TUNE_TABLE = kv["TABLE"]
    det = engine.TileDetector(sd, 3, batch=B, imgsz=(IMG, IMG))
    _lib.TUNE_TABLE = shipped
    for k, v in old.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v
    det.detect(t)
    pred = det.ctx.pred()
    ref = pred if ref is None else ref
    dets[spec] = (det, float(np.abs(pred - ref).max()))
# Interleaved rounds, order reversed every other round; round 0 is warm-up (the GPU settles into its power-capped state: whichever
# spec runs first on a cool chip looks 2-4 % faster); the MEDIAN of the remaining rounds is reported.
ROUNDS = int(os.environ.get("AB_ROUNDS", 9))
times = {s: [] for s in specs}
for r in range(ROUNDS):
    for spec in (specs if r % 2 == 0 else specs[::-1]):
        det = dets[spec][0]
        for _ in range(3):
            det.ctx.forward()
        det.ctx.sync()
        t0 = time.perf_counter()
        for _ in range(25):
            det.ctx.forward()
        det.ctx.sync()
        if r > 0:
            times[spec].append((time.perf_counter() - t0) / 25 * 1e3)
best = {s: float(np.median(times[s])) for s in specs}
for spec in specs:
    print(f"{spec:40s} {best[spec]:7.3f} ms per conv stack (median of {ROUNDS - 1} rounds, min {min(times[spec]):.3f} max {max(times[spec]):.3f})  "
          f"{B / best[spec] * 1e3:8.0f} tiles/s  launches {dets[spec][0].n_launches}  max |pred - first| {dets[spec][1]:.2e}", flush=True)
for d in dets.values():
    d[0].close()
