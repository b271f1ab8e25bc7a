"""TEST INFRASTRUCTURE (oracle) -- CPU restatement of the reference's blank-tile tests.  Not imported by the product.

Follows /root/reference/src/utils.py:325-369:
  is_blank(im)         [REF src/utils.py:325-350]  PIL `im.convert("L").getextrema()`; blank when the extrema are
                       (0,0), (1,1), (255,255) or both >= 250
  is_partly_blank(im)  [REF src/utils.py:353-369]  any row whose mean over (width, channels) is >= 250, or any
                       column whose mean over (height, channels) is >= 250
The measurements are restated with numpy integer arithmetic (PIL's rgb2l is (19595 R + 38470 G + 7471 B + 0x8000) >> 16,
libImaging/Convert.c); tests/golden/blank_ref.npz holds the outputs of the reference's own function source on the same
tiles (tools/make_golden_blank.py) and pins this file.
"""
import numpy as np


def tile_stats(rgb: np.ndarray):
    """rgb: (H, W, 3) uint8 in RGB order -> (l_min, l_max, blank_rows, blank_cols) as the CUDA kernel reports them."""
    a = rgb.astype(np.int64)
    lum = (a[..., 0] * 19595 + a[..., 1] * 38470 + a[..., 2] * 7471 + 0x8000) >> 16
    h, w, _ = rgb.shape
    row_sum = a.sum(axis=(1, 2))
    col_sum = a.sum(axis=(0, 2))
    return int(lum.min()), int(lum.max()), int((row_sum >= 250 * 3 * w).sum()), int((col_sum >= 250 * 3 * h).sum())


def is_blank_from_stats(l_min: int, l_max: int) -> bool:
    ext = (l_min, l_max)
    return ext == (0, 0) or ext == (1, 1) or ext == (255, 255) or (l_min >= 250 and l_max >= 250)


def is_partly_blank_from_stats(blank_rows: int, blank_cols: int) -> bool:
    return blank_rows + blank_cols > 0


def is_blank(rgb: np.ndarray) -> bool:
    s = tile_stats(rgb)
    return is_blank_from_stats(s[0], s[1])


def is_partly_blank(rgb: np.ndarray) -> bool:
    s = tile_stats(rgb)
    return is_partly_blank_from_stats(s[2], s[3])


def windows(raster_h: int, raster_w: int, tile: int):
    """Full windows in the reference's order: x outer, y inner [REF src/load_data/tile_tifs.py:33-47]; (x, y) per window."""
    return [(x, y) for x in range(0, raster_w - tile + 1, tile) for y in range(0, raster_h - tile + 1, tile)]
