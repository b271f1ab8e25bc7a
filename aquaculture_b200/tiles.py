"""Synthetic aerial tiles and tile naming for benchmarks and parity tests.

Real inputs are 1024x1024 JPEG tiles cut by the reference's tiler [REF src/load_data/tile_tifs.py:13,33-37,74]
and named `<LAYER><YYYY>_<bboxind>_<xoff>_<yoff>.jpeg` [REF src/utils.py:372-383]; the consumer splits the
stem on '_' into exactly four fields [REF src/process_yolo/geocode_results.py:89,143].  The corpus is not
in the snapshot, so tiles are uint8 HWC noise from a counter-based generator keyed by (seed, tile index):
any rank can regenerate any tile, and the CPU oracle sees byte-identical input.
"""
from __future__ import annotations

from typing import Iterable, List

import numpy as np


def synthetic_tile(index: int, h: int = 640, w: int = 640, seed: int = 0) -> np.ndarray:
    """uint8 (h, w, 3) BGR-order tile for global tile `index`."""
    rng = np.random.Generator(np.random.Philox(key=(seed << 32) + index))
    return rng.integers(0, 256, (h, w, 3), dtype=np.uint8)


def synthetic_tiles(indices: Iterable[int], h: int = 640, w: int = 640, seed: int = 0) -> np.ndarray:
    idx = list(indices)
    out = np.empty((len(idx), h, w, 3), np.uint8)
    for k, i in enumerate(idx):
        out[k] = synthetic_tile(i, h, w, seed)
    return out


def tile_stem(index: int, year: int = 2020) -> str:
    """A stem the reference's geocoder can parse: 4 '_' fields, field 0 ends with the 4-digit year."""
    bbox_ind, rem = divmod(index, 36)          # 36 tiles per 6144^2 download [REF src/utils.py:17-19]
    x_off, y_off = (rem % 6) * 1024, (rem // 6) * 1024
    return f"ORTHOIMAGERY.ORTHOPHOTOS{year}_{bbox_ind}_{x_off}_{y_off}"


def shard_indices(n_total: int, rank: int, world: int) -> List[int]:
    """Strided shard of the sorted tile list: rank r takes r, r+W, r+2W, ...."""
    return list(range(rank, n_total, world))
