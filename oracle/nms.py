"""ORACLE (test infrastructure): non_max_suppression of the reference's detect path.

Restates upstream `utils/general.py::{non_max_suppression, xywh2xyxy}` [UPSTREAM-RECALLED,
SURVEY.md §8 a-8] as invoked by `yolov5/detect.py` [REF README.md:77] with conf 0.25, IoU 0.45,
max_det 1000, multi_label=False.  The NMS core is the genuine third-party kernel
`torchvision.ops.nms` (CPU; pinned version torchvision==0.15.2 [REF requirements.txt:241], 0.26 here),
so keep-indices are pinned by it; `greedy_nms_numpy` is a second, independent restatement of that
core (SURVEY.md App. C1 semantics) used to cross-check torchvision on small cases.

Deviations from upstream, both deliberate and documented in DESIGN.md:
  * the pre-NMS `argsort(descending=True)` is made stable (ties -> lower row index first), which is
    what torchvision's CPU kernel does on equal scores;
  * upstream's wall-clock time limit (0.5 s + 0.05 s * batch) is not replicated.
"""
from __future__ import annotations

import numpy as np
import torch

MAX_WH = 7680      # upstream: maximum box width/height, used as the per-class coordinate offset
MAX_NMS = 30000    # upstream: maximum number of boxes into torchvision.ops.nms()


def xywh2xyxy(x: torch.Tensor) -> torch.Tensor:
    y = x.clone()
    y[..., 0] = x[..., 0] - x[..., 2] / 2
    y[..., 1] = x[..., 1] - x[..., 3] / 2
    y[..., 2] = x[..., 0] + x[..., 2] / 2
    y[..., 3] = x[..., 1] + x[..., 3] / 2
    return y


def candidates(x: torch.Tensor, conf_thres: float, classes=None):
    """One image (rows, 5+nc) fp32 -> (boxes xyxy, conf, cls, source row index), stably sorted by conf desc."""
    rows = torch.arange(x.shape[0])
    xc = x[:, 4] > conf_thres
    x, rows = x[xc], rows[xc]
    if not x.shape[0]:
        z = torch.zeros(0)
        return torch.zeros(0, 4), z, z, rows
    x = x.clone()
    x[:, 5:] *= x[:, 4:5]
    box = xywh2xyxy(x[:, :4])
    conf, j = x[:, 5:].max(1, keepdim=True)
    keep = conf.view(-1) > conf_thres
    box, conf, j, rows = box[keep], conf[keep].view(-1), j[keep].view(-1).float(), rows[keep]
    if classes is not None:
        m = (j[:, None] == torch.tensor(classes, dtype=torch.float32)).any(1)
        box, conf, j, rows = box[m], conf[m], j[m], rows[m]
    order = torch.argsort(conf, descending=True, stable=True)[:MAX_NMS]
    return box[order], conf[order], j[order], rows[order]


def non_max_suppression(prediction: torch.Tensor, conf_thres=0.25, iou_thres=0.45, classes=None,
                        agnostic=False, max_det=1000, return_rows=False):
    """(B, rows, 5+nc) fp32 -> list of (n,6) [x1,y1,x2,y2,conf,cls] per image (and kept source rows)."""
    import torchvision
    assert 0 <= conf_thres <= 1 and 0 <= iou_thres <= 1
    out, out_rows = [], []
    for x in prediction:
        box, conf, cls, rows = candidates(x.float(), conf_thres, classes)
        if not box.shape[0]:
            out.append(torch.zeros(0, 6))
            out_rows.append(rows)
            continue
        c = cls[:, None] * (0 if agnostic else MAX_WH)
        i = torchvision.ops.nms(box + c, conf, iou_thres)[:max_det]
        out.append(torch.cat((box[i], conf[i, None], cls[i, None]), 1))
        out_rows.append(rows[i])
    return (out, out_rows) if return_rows else out


def greedy_nms_numpy(boxes: np.ndarray, scores: np.ndarray, iou_thres: float) -> np.ndarray:
    """Independent restatement of the torchvision CPU nms core: stable descending order, fp32 IoU
    `inter / (a_i + a_j - inter)`, suppress iff (double)iou > (double)iou_thres."""
    boxes = boxes.astype(np.float32)
    order = np.argsort(-scores.astype(np.float32), kind="stable")
    x1, y1, x2, y2 = (boxes[:, k] for k in range(4))
    areas = ((x2 - x1) * (y2 - y1)).astype(np.float32)
    suppressed = np.zeros(len(boxes), bool)
    keep = []
    for _i in range(len(order)):
        i = order[_i]
        if suppressed[i]:
            continue
        keep.append(i)
        rest = order[_i + 1:]
        xx1 = np.maximum(x1[i], x1[rest])
        yy1 = np.maximum(y1[i], y1[rest])
        xx2 = np.minimum(x2[i], x2[rest])
        yy2 = np.minimum(y2[i], y2[rest])
        w = np.maximum(np.float32(0), (xx2 - xx1).astype(np.float32))
        h = np.maximum(np.float32(0), (yy2 - yy1).astype(np.float32))
        inter = (w * h).astype(np.float32)
        with np.errstate(divide="ignore", invalid="ignore"):
            ovr = (inter / ((areas[i] + areas[rest]).astype(np.float32) - inter).astype(np.float32)).astype(np.float32)
        suppressed[rest[ovr.astype(np.float64) > float(iou_thres)]] = True
    return np.asarray(keep, np.int64)
