"""Builds the shipped tiling table (aquaculture_b200/tuned/b200.txt) on a B200: plans the detector for the batch /
input sizes the CLI, bench.py and the tests use with every feasible tiling of every conv timed on the device
(AQUA_TUNE_ITERS launches each) and collects "signature choice" lines.
  python tools/make_tune_table.py out.txt [B:IMG ...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
out = sys.argv[1]
cfgs = [tuple(int(v) for v in a.split(":")) for a in sys.argv[2:]] or [(64, 640), (16, 1280), (32, 640), (16, 640), (8, 640), (4, 640), (3, 640), (2, 640), (1, 640), (2, 1280), (4, 1280)]
if os.path.exists(out):
    os.remove(out)
os.environ.update(AQUA_AUTOTUNE="1", AQUA_TUNE_CACHE=out, AQUA_TUNE_TABLE="0")
os.environ.setdefault("AQUA_TUNE_LOG", "0")
os.environ.setdefault("AQUA_TUNE_ITERS", "10")
import torch
from aquaculture_b200 import engine, tiles, weights
sd = weights.synthetic_state(3, 0)
for b, img in cfgs:
    for chain in os.environ.get("TUNE_CHAINS", "feed,c3;0;c3;feed").split(";"):  # the default plan and the unchained / partly chained ones (tests, A/B runs)
        os.environ["AQUA_CHAIN"] = chain
        det = engine.TileDetector(sd, 3, batch=b, imgsz=(img, img))
        det.detect(torch.from_numpy(tiles.synthetic_tiles(range(b), img, img)).pin_memory())
        det.close()
    print(f"planned batch {b} at {img}x{img}", flush=True)
lines = sorted(set(open(out).read().splitlines()))
with open(out, "w") as f:
    f.write("# tilings measured on an NVIDIA B200 (tools/make_tune_table.py); one 'signature choice' line per conv shape\n")
    f.write("\n".join(lines) + "\n")
print(f"{len(lines)} entries -> {out}")
