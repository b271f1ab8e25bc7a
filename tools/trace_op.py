"""Per-role timeline of ONE conv on the B200 (instrumented kernel build): where each warp role of a CTA spends a
tile's time.  python tools/trace_op.py B H W c_in c_out k stride [--res] [--mblk N] [--mode M] [--cta 0,1,74]"""
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
ap.add_argument("--cta", default="0,73,147")
ap.add_argument("--tiles", type=int, default=0, help="print per-tile rows for the first CTA")
a = ap.parse_args()
B, H, W, ci, co, k, s = a.shape
g = torch.Generator().manual_seed(0)
w = (torch.randn(co, ci, k, k, generator=g) / math.sqrt(ci * k * k))
b = torch.randn(co, generator=g) * 0.5
ctx = engine.Context(0)
ctx.plan_begin(B, 64, 64, 3)
bs = ctx.buffer(H, W, ci)
bd = ctx.buffer(H // s, W // s, co)
ctx.conv((bs, 0), (bd, 0), w, b, k, s, a.act, (bd, 0) if a.res else None, a.tile_n, a.mblk, a.mode, a.cluster)
ctx.plan_end()
buf = C.create_string_buffer(600)
ctx.call("aq_op_describe", 0, buf, 600)
print(buf.value.decode())
ms = (C.c_float * 1)()
ctx.call("aq_profile_ops", 10, ms, 1)
print(f"untraced: {ms[0]*1e3:.1f} us")
R, T, V = 6, 48, 4
out = np.zeros(148 * R * T * V, np.int64)
grid = ctx.call("aq_trace_op", 0, out.ctypes.data, out.size)
tr = out[:grid * R * T * V].reshape(grid, R, T, V)
names = ["A-prod", "B-prod", "MMA0", "EPI0", "STORE"]
for cta in [int(x) for x in a.cta.split(",") if int(x) < grid]:
    t = tr[cta]
    n = int((t[2, :, 0] != 0).sum())
    if n < 3:
        print(f"cta {cta}: {n} tiles traced"); continue
    lo, hi = 1, n - 1  # steady state: skip first and last tile
    per_tile = (t[2, hi, 0] - t[2, lo, 0]) / (hi - lo)
    print(f"cta {cta}: {n} tiles, steady-state {per_tile:.0f} clk/tile (MMA0 start-to-start)")
    A = t[0, lo:hi]; print(f"  A-prod : busy {np.mean(A[:,3]-A[:,0]):7.0f}  wait a_empty {np.mean(A[:,1]):7.0f}  wait cp.async data {np.mean(A[:,2]):7.0f}")
    Bp = t[1, lo:hi]
    if Bp[:, 0].any(): print(f"  B-prod : busy {np.mean(Bp[:,3]-Bp[:,0]):7.0f}  wait b_empty {np.mean(Bp[:,1]):7.0f}")
    M = t[2, lo:hi]; print(f"  MMA0   : busy {np.mean(M[:,3]-M[:,0]):7.0f}  wait tempty {np.mean(M[:,1]):7.0f}  wait a_full {np.mean(M[:,2]>>32):7.0f}  wait b_full {np.mean(M[:,2]&0xffffffff):7.0f}")
    E = t[3, lo:hi]; print(f"  EPI0   : busy {np.mean(E[:,3]-E[:,0]):7.0f}  wait tfull {np.mean(E[:,1]):7.0f}  wait barrier {np.mean(E[:,2]):7.0f}  work {np.mean(E[:,3]-E[:,0]-E[:,1]-E[:,2]):7.0f}")
    S = t[4, lo:hi]
    if S[:, 0].any():
        lo32 = lambda x: np.mean(x & 0xffffffff); hi32 = lambda x: np.mean(x >> 32)
        print(f"  EPI0 parts/tile: top->ld {lo32(S[:,0]):5.0f} | ld wait {hi32(S[:,0]):5.0f} | tempty arrive {lo32(S[:,1]):5.0f} | process {hi32(S[:,1]):5.0f} | "
              f"prefetch {lo32(S[:,2]):5.0f} | fence+rdwait {hi32(S[:,2]):5.0f} | arrive {lo32(S[:,3]):5.0f}")
    ST = t[5, lo:hi]
    if ST[:, 0].any(): print(f"  STORE  : busy {np.mean(ST[:,3]-ST[:,0]):7.0f}  wait staged group {np.mean(ST[:,1]):7.0f}  wait smem read {np.mean(ST[:,2]):7.0f}")
    if a.tiles and cta == int(a.cta.split(",")[0]):
        t0 = t[2, 0, 0]
        for i in range(min(n, a.tiles)):
            print(f"   tile {i:2d}: A {t[0,i,0]-t0:7d}..{t[0,i,3]-t0:7d} (w {t[0,i,1]:6d} d {t[0,i,2]:6d}) | B {t[1,i,0]-t0:7d}..{t[1,i,3]-t0:7d} (w {t[1,i,1]:6d}) | "
                  f"MMA {t[2,i,0]-t0:7d}..{t[2,i,3]-t0:7d} (te {t[2,i,1]:6d} a {t[2,i,2]>>32:6d} b {t[2,i,2]&0xffffffff:6d}) | "
                  f"EPI {t[3,i,0]-t0:7d}..{t[3,i,3]-t0:7d} (tf {t[3,i,1]:6d} se {t[3,i,2]:6d})")
ctx.close()
