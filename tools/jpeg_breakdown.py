"""Where the time of `.jpeg bytes -> detections` goes (bench.py's jpeg_ingest leg), per 64-tile batch of 640^2 tiles:
host prepare (aq_jpeg_scan_prepare_batch), submit call, GPU decode alone (CUDA events around aq_jpeg_decode_device), the
detector alone, and the pipelined loop.   python tools/jpeg_breakdown.py [--size 640]"""
import argparse, ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cv2, numpy as np, torch
from aquaculture_b200 import engine, tiles, weights
ap = argparse.ArgumentParser()
ap.add_argument("--size", type=int, default=640)
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
cv2.setNumThreads(1)
rng = np.random.default_rng(0)
blobs = []
for i in range(a.batch):
    base = rng.integers(0, 256, (a.size // 16 + 2, a.size // 16 + 2, 3), dtype=np.uint8)
    img = cv2.resize(base, (a.size, a.size), interpolation=cv2.INTER_CUBIC).astype(np.int32) + rng.integers(-12, 12, (a.size, a.size, 3))
    blobs.append(cv2.imencode(".jpg", np.clip(img, 0, 255).astype(np.uint8), [cv2.IMWRITE_JPEG_QUALITY, 75])[1])
threads = len(os.sched_getaffinity(0))
det = engine.TileDetector(weights.synthetic_state(3, 0), 3, batch=a.batch, imgsz=(640, 640))
def T(fn, reps=a.reps):
    fn(); det.ctx.sync()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    det.ctx.sync()
    return (time.perf_counter() - t0) / reps * 1e3
prep = lambda: engine.prepare_jpeg_batch(det, blobs, n_threads=threads)
print(f"host prepare ({threads} threads):      {T(prep):7.3f} ms per batch")
(sc, off, ln, tab, q, stt), tot, i0 = prep()
out = torch.empty(a.batch, i0.height, i0.width, 3, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
dec = lambda: det.ctx.call("aq_jpeg_decode_device", C.c_void_p(sc.ctypes.data), C.c_void_p(off.ctypes.data), C.c_void_p(ln.ctypes.data), int(tot),
                           C.c_void_p(tab.ctypes.data), C.c_void_p(q.ctypes.data), a.batch, i0.width, i0.height, i0.ncomp, i0.hs, 1,
                           C.c_void_p(out.data_ptr()), C.c_void_p(stt.ctypes.data), None)
print(f"GPU decode alone (H2D + kernels):     {T(dec):7.3f} ms per batch  ({tot / 1e6:.2f} MB of scans)")
dev = torch.from_numpy(tiles.synthetic_tiles(range(a.batch), a.size, a.size)).cuda()
print(f"detector alone (resident pixels):     {T(lambda: det.detect_wait(det.detect_submit(dev, bgr=True, src_is_device=True))):7.3f} ms per batch")
def sub():
    (sc, off, ln, tab, q, stt), tot, i0 = engine.prepare_jpeg_batch(det, blobs, n_threads=threads)
    return det.detect_submit_jpeg_scans(sc, off, ln, tot, tab, q, stt, i0.width, i0.height, i0.ncomp, i0.hs)
print(f"decode + detect, one batch at a time: {T(lambda: det.detect_wait(sub())):7.3f} ms per batch")
t_sub = []
def loop(n):
    prev = None
    for _ in range(n):
        t0 = time.perf_counter(); tk = sub(); t_sub.append(time.perf_counter() - t0)
        if prev is not None:
            det.detect_wait(prev)
        prev = tk
    det.detect_wait(prev)
loop(3)
t_sub.clear()
t0 = time.perf_counter(); loop(a.reps); dt = time.perf_counter() - t0
print(f"pipelined (two batches in flight):    {dt / a.reps * 1e3:7.3f} ms per batch = {a.batch * a.reps / dt:.0f} tiles/s; host time inside prepare+submit {np.mean(t_sub) * 1e3:.3f} ms")
det.close()

# GPU time of the two halves INSIDE the pipelined loop (CUDA events on the launching stream)
det = engine.TileDetector(weights.synthetic_state(3, 0), 3, batch=a.batch, imgsz=(640, 640))
st = torch.cuda.Stream()
sarg = C.c_void_p(st.cuda_stream)
ev = []
def sub2():
    (sc, off, ln, tab, q, stt), tot, i0 = engine.prepare_jpeg_batch(det, blobs, n_threads=threads)
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    with torch.cuda.stream(st):
        e[0].record()
        key = (i0.height, i0.width)
        if det._jpeg_out.get(key) is None:
            det._jpeg_out[key] = torch.empty(det.batch, i0.height, i0.width, 3, dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
        out = det._jpeg_out[key][:a.batch]
        det.ctx.call("aq_jpeg_decode_device", C.c_void_p(sc.ctypes.data), C.c_void_p(off.ctypes.data), C.c_void_p(ln.ctypes.data), int(tot),
                     C.c_void_p(tab.ctypes.data), C.c_void_p(q.ctypes.data), a.batch, i0.width, i0.height, i0.ncomp, i0.hs, 1,
                     C.c_void_p(out.data_ptr()), C.c_void_p(stt.ctypes.data), sarg)
        e[1].record()
        tk = det.detect_submit(out, bgr=True, src_is_device=True, stream=sarg)
        e[2].record()
    ev.append(e)
    return tk
prev = None
for i in range(a.reps + 3):
    tk = sub2()
    if prev is not None:
        det.detect_wait(prev)
    prev = tk
det.detect_wait(prev)
torch.cuda.synchronize()
d = np.array([[e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])] for e in ev[3:]])
gap = np.array([ev[i][2].elapsed_time(ev[i + 1][0]) for i in range(3, len(ev) - 1)])
print(f"in the pipelined loop, GPU time per batch: decode {d[:, 0].mean():.3f} ms, detect {d[:, 1].mean():.3f} ms, idle between batches {gap.mean():.3f} ms")
det.close()
