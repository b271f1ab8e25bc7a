"""ORACLE (test infrastructure): box rescale + label-file text of the reference's detect path.

Restates upstream `utils/general.py::{scale_boxes, clip_boxes, xyxy2xywh}` and the `--save-txt`
block of `detect.py::run` [UPSTREAM-RECALLED, SURVEY.md §8 a-9].  The grammar is pinned by the
consumer, `np.loadtxt` + `bbox[0..5]` in [REF src/process_yolo/geocode_results.py:140-172]:
one line per detection `cls cx cy w h conf`, '%g' each, normalised by the ORIGINAL image size,
ascending confidence (upstream iterates `reversed(det)`), no file for zero detections.
"""
from __future__ import annotations

import os
from typing import List

import torch


def scale_boxes(img1_shape, boxes: torch.Tensor, img0_shape) -> torch.Tensor:
    """Rescale xyxy boxes from the letterboxed shape img1 (h, w) to the original img0 (h, w); clip."""
    gain = min(img1_shape[0] / img0_shape[0], img1_shape[1] / img0_shape[1])
    pad = (img1_shape[1] - img0_shape[1] * gain) / 2, (img1_shape[0] - img0_shape[0] * gain) / 2
    boxes = boxes.clone()
    boxes[..., [0, 2]] -= pad[0]
    boxes[..., [1, 3]] -= pad[1]
    boxes[..., :4] /= gain
    boxes[..., 0].clamp_(0, img0_shape[1])
    boxes[..., 1].clamp_(0, img0_shape[0])
    boxes[..., 2].clamp_(0, img0_shape[1])
    boxes[..., 3].clamp_(0, img0_shape[0])
    return boxes


def xyxy2xywh(x: torch.Tensor) -> torch.Tensor:
    y = x.clone()
    y[..., 0] = (x[..., 0] + x[..., 2]) / 2
    y[..., 1] = (x[..., 1] + x[..., 3]) / 2
    y[..., 2] = x[..., 2] - x[..., 0]
    y[..., 3] = x[..., 3] - x[..., 1]
    return y


def label_lines(det: torch.Tensor, img1_shape, img0_shape, save_conf=True) -> List[str]:
    """det (n,6) fp32 in letterboxed pixels, descending conf -> text lines in file order."""
    if not len(det):
        return []
    det = det.clone().float()
    det[:, :4] = scale_boxes(img1_shape, det[:, :4], img0_shape).round()
    gn = torch.tensor(img0_shape)[[1, 0, 1, 0]]
    lines = []
    for *xyxy, conf, cls in reversed(det):
        xywh = (xyxy2xywh(torch.tensor(xyxy).view(1, 4)) / gn).view(-1).tolist()
        line = (cls, *xywh, conf) if save_conf else (cls, *xywh)
        lines.append(("%g " * len(line)).rstrip() % line)
    return lines


def write_labels(label_dir: str, stem: str, lines: List[str]) -> bool:
    """Append-mode write as upstream does; no file when there is nothing to write."""
    if not lines:
        return False
    with open(os.path.join(label_dir, stem + ".txt"), "a") as f:
        f.write("".join(l + "\n" for l in lines))
    return True
