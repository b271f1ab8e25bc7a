"""ORACLE (test infrastructure) and timed CPU baseline: upstream `detect.py::run`, restated.

PARITY UNPINNED end to end (see oracle/yolov5m.py).  Follows the per-tile loop of
`yolov5/detect.py` [REF README.md:77; UPSTREAM-RECALLED SURVEY.md §3.1]: batch 1, fp32,
letterbox(auto=True) -> CHW RGB /255 -> model -> non_max_suppression -> scale_boxes().round()
-> '%g' label lines appended to labels/<stem>.txt.  Only tests/, __graft_entry__.smoke() and
bench.py's CPU-baseline / --impl reference legs may import this module.
"""
from __future__ import annotations

import os
import time
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import labels as L
from . import letterbox as LB
from . import nms as N
from . import yolov5m


class DetectRef:
    def __init__(self, state_dict: Dict[str, torch.Tensor], nc: int = 3, imgsz=(640, 640),
                 conf_thres=0.25, iou_thres=0.45, max_det=1000, classes=None, agnostic=False,
                 threads: Optional[int] = None):
        if threads:
            torch.set_num_threads(threads)
        self.model = yolov5m.build(state_dict, nc)
        self.imgsz = tuple(imgsz)
        self.kw = dict(conf_thres=conf_thres, iou_thres=iou_thres, classes=classes, agnostic=agnostic,
                       max_det=max_det)
        self.dt = [0.0, 0.0, 0.0]
        self.seen = 0

    @torch.no_grad()
    def predict(self, im_bgr: np.ndarray):
        """One HWC BGR uint8 tile -> (det (n,6) in letterboxed pixels, letterboxed (h,w), raw pred)."""
        t0 = time.perf_counter()
        lb = LB.letterbox(im_bgr, self.imgsz, auto=True)
        x = torch.from_numpy(LB.to_tensor(lb))[None]
        t1 = time.perf_counter()
        pred = self.model(x)
        t2 = time.perf_counter()
        det = N.non_max_suppression(pred, **self.kw)[0]
        t3 = time.perf_counter()
        self.dt[0] += t1 - t0
        self.dt[1] += t2 - t1
        self.dt[2] += t3 - t2
        self.seen += 1
        return det, tuple(x.shape[2:]), pred

    def run(self, tiles: Sequence[np.ndarray], stems: Sequence[str], label_dir: Optional[str],
            save_conf=True) -> List[List[str]]:
        all_lines = []
        for im, stem in zip(tiles, stems):
            det, shape1, _ = self.predict(im)
            lines = L.label_lines(det, shape1, im.shape[:2], save_conf)
            if label_dir is not None:
                L.write_labels(label_dir, stem, lines)
            all_lines.append(lines)
        return all_lines


def time_cpu_baseline(state_dict, nc, tiles: Sequence[np.ndarray], warmup: int = 2,
                      threads: Optional[int] = None, **kw) -> dict:
    """tiles/s of the restated detect.py on this host's cores (batch 1, fp32), incl. label text."""
    threads = threads or len(os.sched_getaffinity(0))
    ref = DetectRef(state_dict, nc, threads=threads, **kw)
    for im in tiles[:warmup]:
        ref.predict(im)
    ref.dt, ref.seen = [0.0, 0.0, 0.0], 0
    t0 = time.perf_counter()
    ref.run(tiles[warmup:], [f"t{i}" for i in range(len(tiles) - warmup)], None)
    wall = time.perf_counter() - t0
    n = len(tiles) - warmup
    return dict(tiles_per_s=n / wall, n=n, wall_s=wall, cores=threads,
                stage_ms=[1e3 * d / max(n, 1) for d in ref.dt])
