"""Golden vectors for the blank-tile tests: runs the reference's OWN is_blank / is_partly_blank source
(/root/reference/src/utils.py, extracted with ast so that its geopandas imports are not needed) on seeded tiles.
Run in the build container only (needs /root/reference): python tools/make_golden_blank.py -> tests/golden/blank_ref.npz"""
import ast
import os
from io import BytesIO

import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = open("/root/reference/src/utils.py").read()
mod = ast.parse(src)
wanted = [n for n in mod.body if isinstance(n, ast.FunctionDef) and n.name in ("is_blank", "is_partly_blank")]
ns = {"Image": Image, "BytesIO": BytesIO, "np": np}
exec(compile(ast.Module(body=wanted, type_ignores=[]), "reference:src/utils.py", "exec"), ns)

rng = np.random.default_rng(5)
tiles = []
H, W = 96, 128


def t(fill=None):
    return rng.integers(0, 256, (H, W, 3), dtype=np.uint8) if fill is None else np.full((H, W, 3), fill, np.uint8)


tiles += [t(), t(0), t(1), t(255), t(250), t(249), t(252)]
a = t(); a[10:13] = 255; tiles.append(a)                       # blank rows
a = t(); a[:, 40:44] = 252; tiles.append(a)                    # blank columns
a = t(); a[5] = 250; a[:, 7] = 249; tiles.append(a)            # row exactly 250, column just below
a = t(253); a[0, 0] = (0, 0, 0); tiles.append(a)               # nearly white with one black pixel
a = t(); a[20] = (255, 255, 239); tiles.append(a)              # row mean 249.67
a = t(); a[20] = (255, 255, 240); tiles.append(a)              # row mean exactly 250
a = rng.integers(250, 256, (H, W, 3), dtype=np.uint8); tiles.append(a)   # all luminance >= 250
a = rng.integers(0, 2, (H, W, 3), dtype=np.uint8); tiles.append(a)       # luminance 0/1 mix
a = t(1); tiles.append(a)
tiles = np.stack(tiles)
blank = np.array([bool(ns["is_blank"](im=Image.fromarray(x))) for x in tiles])
partly = np.array([bool(ns["is_partly_blank"](Image.fromarray(x))) for x in tiles])
extrema = np.array([Image.fromarray(x).convert("L").getextrema() for x in tiles], dtype=np.int32)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "blank_ref.npz"), tiles=tiles, is_blank=blank, is_partly_blank=partly,
                    extrema=extrema)
print("tiles", tiles.shape, "blank", blank.astype(int), "partly", partly.astype(int))
