"""Throughput of K detector contexts on ONE GPU, each on its own stream with its own activation buffers, batches submitted
round-robin: one context's launch / drain bubbles and tail waves are filled by the other's kernels (one 672-thread CTA per
SM: a kernel of stream B takes the SMs stream A's CTAs leave).  Prints tiles/s for K = 1, 2, 3 (device-resident tiles,
preprocess + conv stack + NMS + D2H of the detections).   python tools/two_stream_probe.py [--batches 60]"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from aquaculture_b200 import engine, tiles, weights
ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--img", type=int, default=640)
ap.add_argument("--batches", type=int, default=60)
ap.add_argument("--k", type=int, nargs="+", default=[1, 2, 3])
a = ap.parse_args()
sd = weights.synthetic_state(3, 0)
sets = [torch.from_numpy(tiles.synthetic_tiles(range(s * a.batch, (s + 1) * a.batch), a.img, a.img)).cuda() for s in range(4)]
torch.cuda.synchronize()
ref = None
for K in a.k:
    dets = [engine.TileDetector(sd, 3, batch=a.batch, imgsz=(a.img, a.img)) for _ in range(K)]
    def run(n):
        pend = []
        out = []
        for i in range(n):
            d = dets[i % K]
            pend.append((d, d.detect_submit(sets[i % 4], bgr=True, src_is_device=True)))
            if len(pend) > 2 * K - 1:  # every context keeps two batches in flight
                dd, t = pend.pop(0)
                out.append(dd.detect_wait(t).counts.copy())
        for dd, t in pend:
            out.append(dd.detect_wait(t).counts.copy())
        return out
    run(3 * K)
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        counts = run(a.batches)
        best = min(best, time.perf_counter() - t0)
    c = np.concatenate(counts)
    if ref is None:
        ref = c
    print(f"contexts {K}: {a.batches * a.batch / best:9.0f} tiles/s  ({best / a.batches * 1e3:.3f} ms per batch)  pdl={os.environ.get('AQUA_PDL', '1')}  "
          f"same detection counts as K=1: {bool(np.array_equal(c, ref))}", flush=True)
    for d in dets:
        d.close()
