"""Generates tests/golden/*.npz from the genuine third-party kernels the reference path dispatches to
(torchvision.ops.nms CPU, cv2.resize / copyMakeBorder) — the only pieces of the reference's arithmetic
that can be executed in this container.  Run: python tools/make_golden.py"""
import os
import sys

import cv2
import numpy as np
import torch
import torchvision

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")


def nms_cases():
    g = torch.Generator().manual_seed(1234)
    cases = {}
    for name, n, scale, quant in [("random300", 300, 60.0, None), ("ties200", 200, 80.0, 16), ("dense500", 500, 200.0, None)]:
        xy = torch.rand(n, 2, generator=g) * 640
        wh = torch.rand(n, 2, generator=g) * scale + 2
        boxes = torch.cat((xy - wh / 2, xy + wh / 2), 1)
        scores = torch.rand(n, generator=g)
        if quant:
            scores = (scores * quant).round() / quant
        for thr in (0.45, 0.1):
            keep = torchvision.ops.nms(boxes, scores, thr)
            cases[f"{name}_{thr}_boxes"] = boxes.numpy()
            cases[f"{name}_{thr}_scores"] = scores.numpy()
            cases[f"{name}_{thr}_keep"] = keep.numpy()
    # threshold-is-a-double-compare vectors
    b = torch.tensor([[0, 0, 1, 5.5], [0, 4.5, 1, 10]])
    s = torch.tensor([0.9, 0.8])
    cases["edge_0.1_boxes"], cases["edge_0.1_scores"] = b.numpy(), s.numpy()
    cases["edge_0.1_keep"] = torchvision.ops.nms(b, s, 0.1).numpy()
    np.savez_compressed(os.path.join(OUT, "nms_torchvision.npz"), **cases)
    print("nms cases:", len(cases) // 3, "torchvision", torchvision.__version__)


def resize_cases():
    rng = np.random.default_rng(7)
    cases = {}
    for name, (h, w), (nh, nw) in [("1024to640", (96, 1024), (60, 640)), ("down_odd", (75, 133), (48, 85)),
                                   ("up", (30, 50), (77, 128)), ("half", (64, 64), (32, 32))]:
        im = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        cases[name + "_src"] = im
        cases[name + "_dst"] = cv2.resize(im, (nw, nh), interpolation=cv2.INTER_LINEAR)
    np.savez_compressed(os.path.join(OUT, "resize_cv2.npz"), **cases)
    print("resize cases:", len(cases) // 2, "cv2", cv2.__version__)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    nms_cases()
    resize_cases()
