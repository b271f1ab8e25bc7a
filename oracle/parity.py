"""ORACLE (test infrastructure): comparison helpers shared by tests/ and bench.py's `parity` leg.

Detections of the CUDA path against the oracle's (oracle/detect_ref.py, PARITY UNPINNED end to end -- see
oracle/yolov5m.py): IoU matching per class, and a proof obligation for every detection that exists on one side only.
A one-sided detection is accepted only when it is a THRESHOLD FLIP of upstream's own rules
[UPSTREAM-RECALLED utils/general.py::non_max_suppression]:
  * confidence flip: its confidence lies within `conf_tol` of conf_thres (obj * cls > conf_thres decides candidacy);
  * NMS flip: a higher-scored detection of the same class overlaps it with an IoU within `iou_tol` of iou_thres
    (torchvision.ops.nms suppresses at IoU > iou_thres), i.e. fp16-level box differences moved the pair across it.
Anything else counts as `unexplained` and fails the caller.  Never imported by the product path.
"""
from __future__ import annotations

import numpy as np
import torch
from torchvision.ops import box_iou


def match_detections(got: torch.Tensor, ref: torch.Tensor, conf_thres=0.25, iou_thres=0.45, conf_tol=5e-3, iou_tol=2e-2):
    """got / ref: (n, 6) x1 y1 x2 y2 conf cls.  Returns dict(box_err, conf_err, matched, flips, unexplained)."""
    got, ref = got.float(), ref.float()
    out = dict(box_err=0.0, conf_err=0.0, matched=0, flips=0, unexplained=0)
    got_ok = torch.zeros(len(got), dtype=torch.bool)
    ref_hit = torch.zeros(len(ref), dtype=torch.bool)
    if len(got) and len(ref):
        iou = box_iou(got[:, :4], ref[:, :4])
        iou[got[:, 5][:, None] != ref[:, 5][None, :]] = 0
        pair = torch.full((len(got),), -1, dtype=torch.long)
        for i in torch.argsort(got[:, 4], descending=True).tolist():  # one-to-one, most confident first
            row = iou[i].clone()
            row[ref_hit] = 0
            v, j = row.max(0)
            if v > 0.5:
                pair[i] = j
                ref_hit[j] = True
                got_ok[i] = True
        if got_ok.any():
            j = pair[got_ok]
            diag = (ref[j, 2:4] - ref[j, 0:2]).norm(dim=1).clamp(min=1.0)
            out["box_err"] = float(((got[got_ok, :4] - ref[j, :4]).abs().max(1).values / diag).max())
            out["conf_err"] = float((got[got_ok, 4] - ref[j, 4]).abs().max())
            out["matched"] = int(got_ok.sum())
    lone = [(got[i], got, ref) for i in torch.nonzero(~got_ok).flatten()] + [(ref[i], ref, got) for i in torch.nonzero(~ref_hit).flatten()]
    for d, own, other in lone:
        if abs(float(d[4]) - conf_thres) < conf_tol:
            out["flips"] += 1
            continue
        pool = torch.cat((own, other), 0)
        pool = pool[(pool[:, 5] == d[5]) & (pool[:, 4] > d[4] - conf_tol)]
        flip = False
        if len(pool):
            v = box_iou(d[None, :4], pool[:, :4])[0]
            flip = bool(((v > iou_thres - iou_tol) & (v < iou_thres + iou_tol)).any())
        out["flips" if flip else "unexplained"] += 1
    return out


def activation_close(got: np.ndarray, ref: np.ndarray, atol=2e-2, rtol=1e-2):
    """fp16-scaled closeness of an intermediate activation: |got - ref| <= atol + rtol * |ref| everywhere.
    Returns (ok, worst excess ratio, max abs error)."""
    d = np.abs(got.astype(np.float32) - ref.astype(np.float32))
    lim = atol + rtol * np.abs(ref.astype(np.float32))
    return bool((d <= lim).all()), float((d / lim).max()), float(d.max())


def oracle_activations(model, x: torch.Tensor, layers):
    """Forward the oracle model on x (NCHW fp32) capturing the outputs of top-level layers `layers` as NHWC fp32 arrays.
    Returns (pred, {layer: array})."""
    acts, hooks = {}, []
    for i in layers:
        hooks.append(model.model[i].register_forward_hook(
            lambda m, inp, out, i=i: acts.__setitem__(i, out.detach().permute(0, 2, 3, 1).contiguous().numpy())))
    with torch.no_grad():
        pred = model(x)
    for h in hooks:
        h.remove()
    return pred, acts
