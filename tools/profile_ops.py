"""Per-op timing table of the planned forward (CUDA events, each op alone, warm): achieved TFLOP/s and GB/s
against the measured peaks.  python tools/profile_ops.py [--batch 64] [--img 640] [--iters 10]"""
import argparse, ctypes as C, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from aquaculture_b200 import engine, tiles, weights

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--img", type=int, default=640)
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--out", default="")
ap.add_argument("--insitu", action="store_true", help="events between the ops of whole forward passes")
a = ap.parse_args()
sd = weights.synthetic_state(3, 0)
det = engine.TileDetector(sd, 3, batch=a.batch, imgsz=(a.img, a.img))
t = torch.from_numpy(tiles.synthetic_tiles(range(a.batch), a.img, a.img)).pin_memory()
det.detect(t)
n = det.n_launches
ms = (C.c_float * n)()
det.ctx.call("aq_profile_ops_insitu" if a.insitu else "aq_profile_ops", a.iters, ms, n)
rows, tot_ms, tot_flop, tot_bytes = [], 0.0, 0.0, 0.0
buf = C.create_string_buffer(512)
for i in range(n):
    det.ctx.call("aq_op_describe", i, buf, 512)
    d = buf.value.decode()
    flop = float(re.search(r"flop=([0-9.e+]+)", d).group(1)); byts = float(re.search(r"bytes=([0-9.e+]+)", d).group(1))
    tot_ms += ms[i]; tot_flop += flop; tot_bytes += byts
    rows.append((i, ms[i], flop / ms[i] / 1e9, byts / ms[i] / 1e6, d))
lines = [f"{i:3d} {m*1e3:8.1f}us {tf:7.1f}TF/s {gb:7.0f}GB/s  {d}" for i, m, tf, gb, d in rows]
lines.append(f"TOTAL {tot_ms:.3f} ms  {tot_flop/tot_ms/1e9:.1f} TFLOP/s  {tot_bytes/tot_ms/1e6:.0f} GB/s (unfused algorithmic bytes {tot_bytes/1e6/a.batch:.1f} MB/tile)  "
             f"-> {a.batch/tot_ms*1e3:.0f} tiles/s conv-stack only")
print("\n".join(lines))
if a.out:
    open(a.out, "w").write("\n".join(lines) + "\n")
det.close()
