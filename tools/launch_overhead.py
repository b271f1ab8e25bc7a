"""Fixed cost of one conv launch: every planned conv restricted to 1, 2, 4, 8 waves of tiles, launched back to back
(aq_profile_op_waves).  Prints per op: time per launch at each wave count, the per-wave slope and the intercept.
  python tools/launch_overhead.py [--batch 64] [--img 640] [--iters 30]"""
import argparse, ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from aquaculture_b200 import engine, tiles, weights

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--img", type=int, default=640)
ap.add_argument("--iters", type=int, default=30)
a = ap.parse_args()
det = engine.TileDetector(weights.synthetic_state(3, 0), 3, batch=a.batch, imgsz=(a.img, a.img))
det.detect(torch.from_numpy(tiles.synthetic_tiles(range(a.batch), a.img, a.img)).pin_memory())
buf = C.create_string_buffer(512)
W = (1, 2, 4, 8)
tot_fix = 0.0
for i in range(det.n_launches):
    det.ctx.call("aq_op_describe", i, buf, 512)
    d = buf.value.decode()
    if not d.startswith(("conv", "detect")):
        continue
    ms = C.c_float()
    us, tl = [], []
    for w in W:
        tl.append(det.ctx.call("aq_profile_op_waves", i, w, a.iters, C.byref(ms)))
        us.append(ms.value * 1e3)
    full = C.c_float()
    n_full = det.ctx.call("aq_profile_op_waves", i, 1 << 20, a.iters, C.byref(full))
    k = [j for j in range(len(W)) if tl[j] == W[j] * 148]
    if len(k) >= 2:
        slope, icpt = np.polyfit([W[j] for j in k], [us[j] for j in k], 1)
    else:
        slope, icpt = float("nan"), float("nan")
    tot_fix += 0 if np.isnan(icpt) else icpt
    print(f"{i:3d} " + " ".join(f"{u:7.1f}" for u in us) + f"  full {full.value*1e3:7.1f}us ({n_full} tiles)  wave {slope:6.2f}us  fixed {icpt:6.2f}us | {d[:70]}", flush=True)
print(f"sum of fixed costs: {tot_fix:.1f} us")
det.close()
