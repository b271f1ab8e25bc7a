"""Plan ONE conv of a given shape on the B200, time it (CUDA events inside the library) and print its plan.
Meant for ncu captures of a single layer:  python tools/one_op.py 64 320 320 16 48 3 1 [--res] [--mblk N] [--mode M]"""
import argparse, ctypes as C, math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from aquaculture_b200 import engine

ap = argparse.ArgumentParser()
ap.add_argument("shape", type=int, nargs=7, help="B H W c_in c_out k stride")
ap.add_argument("--res", action="store_true")
ap.add_argument("--mblk", type=int, default=0)
ap.add_argument("--mode", type=int, default=0)
ap.add_argument("--tile-n", type=int, default=0)
ap.add_argument("--cluster", type=int, default=0)
ap.add_argument("--act", type=int, default=2)
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--src-ct", type=int, default=0, help="channels of the source buffer (the conv reads the slice [src_off, src_off + c_in))")
ap.add_argument("--src-off", type=int, default=0)
ap.add_argument("--dst-ct", type=int, default=0)
ap.add_argument("--dst-off", type=int, default=0)
a = ap.parse_args()
B, H, W, ci, co, k, s = a.shape
g = torch.Generator().manual_seed(0)
w = (torch.randn(co, ci, k, k, generator=g) / math.sqrt(ci * k * k))
b = torch.randn(co, generator=g) * 0.5
ctx = engine.Context(0)
ctx.plan_begin(B, 64, 64, 3)
sct, dct = a.src_ct or ci, a.dst_ct or co
bs = ctx.buffer(H, W, sct)
bd = ctx.buffer(H // s, W // s, dct)
ctx.conv((bs, a.src_off), (bd, a.dst_off), w, b, k, s, a.act, (bd, a.dst_off) if a.res else None, a.tile_n, a.mblk, a.mode, a.cluster)
ctx.plan_end()
x = torch.randn(1, H, W, sct, generator=g).half().numpy()
ctx.upload(bs, np.ascontiguousarray(np.broadcast_to(x, (B, H, W, sct))))
ms = (C.c_float * 1)()
ctx.call("aq_profile_ops", a.iters, ms, 1)
buf = C.create_string_buffer(600)
ctx.call("aq_op_describe", 0, buf, 600)
d = buf.value.decode()
flop = 2.0 * B * (H // s) * (W // s) * co * ci * k * k
print(f"{ms[0]*1e3:9.1f} us  {flop/ms[0]/1e9:7.1f} TF/s  {d}")
ctx.close()
