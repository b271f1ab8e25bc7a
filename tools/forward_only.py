"""Smallest program that runs the benchmarked conv stack: plans the batch-64 640x640 detector (shipped tilings), one whole
detect() and `--warm` forward passes as warm-up, then `--passes` forward passes.  For ncu: every forward pass launches 75
conv_igemm / sppf kernels (-k regex:'conv_igemm|sppf' -s 75*(1+warm) -c 75)."""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from aquaculture_b200 import engine, tiles, weights
ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--img", type=int, default=640)
ap.add_argument("--warm", type=int, default=2)
ap.add_argument("--passes", type=int, default=1)
a = ap.parse_args()
det = engine.TileDetector(weights.synthetic_state(3, 0), 3, batch=a.batch, imgsz=(a.img, a.img))
r = det.detect(torch.from_numpy(tiles.synthetic_tiles(range(a.batch), a.img, a.img)).pin_memory())
for _ in range(a.warm + a.passes):
    det.ctx.forward()
det.ctx.sync()
print(f"forward_only: {det.n_launches} launches per pass, {int(r.counts.sum())} detections, tune misses {det.ctx.call('aq_tune_misses')}")
det.close()
