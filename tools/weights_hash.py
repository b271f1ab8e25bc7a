"""sha256 of the seeded stand-in weights (weights.synthetic_state) and of 4 synthetic tiles: both are generated on the host, so two
machines agree on label files of `--weights synthetic` runs only if these hashes agree.  python tools/weights_hash.py"""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from aquaculture_b200 import tiles, weights
sd = weights.synthetic_state(3, 0)
h = hashlib.sha256()
for k in sorted(sd):
    h.update(k.encode()); h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
t = hashlib.sha256(tiles.synthetic_tiles(range(4), 640, 640).tobytes()).hexdigest()[:16]
print("weights", h.hexdigest()[:16], "tiles", t, "| torch", torch.__version__, "cpu", torch.backends.cpu.get_cpu_capability())
