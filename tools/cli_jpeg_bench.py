"""detect.py on a directory of .jpeg tiles, end to end (file reads, decode, detect, label files), per --jpeg-decode mode:
python tools/cli_jpeg_bench.py [--tiles 1024] [--size 1024].  Writes the tiles to a temporary directory first (synthetic
scenes, quality 75, 4:2:0 -- what gdal.Translate -of JPEG produces), then runs the CLI in-process once per mode and prints
one JSON line: tiles/s per mode (wall clock of the whole CLI call, model planning included, and of the steady state the
CLI reports), label-file counts and whether all modes wrote identical label files."""
import argparse
import contextlib
import io
import json
import os
import re
import shutil
import sys
import tempfile
import time
from concurrent.futures import ThreadPoolExecutor

import cv2
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aquaculture_b200 import detect_cli as cli  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--tiles", type=int, default=1024)
ap.add_argument("--size", type=int, default=1024)
ap.add_argument("--warm", type=int, default=64, help="tiles of the untimed warm-up call (CUDA init, module load)")
a = ap.parse_args()
# tilings come from the shipped table (identical in every run): label files are comparable bit for bit
root = tempfile.mkdtemp(prefix="aq_jpegs_")
src = os.path.join(root, "jpegs")
os.makedirs(src)


def write(i):
    rng = np.random.default_rng(i)
    s = a.size
    base = rng.integers(0, 256, (s // 16 + 2, s // 16 + 2, 3), dtype=np.uint8)
    img = cv2.resize(base, (s, s), interpolation=cv2.INTER_CUBIC).astype(np.int32) + rng.integers(-12, 12, (s, s, 3))
    cv2.imwrite(os.path.join(src, f"ORTHO2020_{i}_0_0.jpeg"), np.clip(img, 0, 255).astype(np.uint8), [cv2.IMWRITE_JPEG_QUALITY, 75])


with ThreadPoolExecutor(len(os.sched_getaffinity(0))) as ex:
    list(ex.map(write, range(a.tiles)))
res = {"tiles": a.tiles, "size": a.size, "host_threads": len(os.sched_getaffinity(0)),
       "jpeg_bytes_per_tile": int(np.mean([os.path.getsize(os.path.join(src, f)) for f in os.listdir(src)]))}
labels = {}
if a.warm > 0:  # untimed: CUDA context, module load, first plan
    wsrc = os.path.join(root, "warm")
    os.makedirs(wsrc)
    for f in sorted(os.listdir(src))[:a.warm]:
        os.symlink(os.path.join(src, f), os.path.join(wsrc, f))
    with contextlib.redirect_stdout(io.StringIO()):
        cli.main(["--weights", "synthetic", "--source", wsrc, "--nosave", "--save-txt", "--save-conf", "--project", os.path.join(root, "runs"),
                  "--name", "warm"])
for mode in ("device", "hybrid", "host"):
    argv = ["--weights", "synthetic", "--source", src, "--nosave", "--save-txt", "--save-conf", "--project", os.path.join(root, "runs"),
            "--name", mode, "--jpeg-decode", mode, "--io-threads", str(res["host_threads"])]
    buf = io.StringIO()
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(buf):
        rc = cli.main(argv)
    wall = time.perf_counter() - t0
    m = re.search(r"([0-9.]+) tiles/s", buf.getvalue())
    ld = os.path.join(root, "runs", mode, "labels")
    labels[mode] = {f: open(os.path.join(ld, f)).read() for f in sorted(os.listdir(ld))}
    res[mode] = {"rc": rc, "tiles_per_s_whole_call": round(a.tiles / wall, 1), "tiles_per_s_reported": float(m.group(1)) if m else None,
                 "label_files": len(labels[mode])}
res["identical_label_files"] = labels["device"] == labels["hybrid"] == labels["host"]
shutil.rmtree(root, ignore_errors=True)
print(json.dumps(res))
