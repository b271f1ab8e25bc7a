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


def compare_image(pred_got, dets_got, pred_ref, conf_thres=0.25, iou_thres=0.45, max_det=1000, conf_tol=5e-3, iou_tol=2e-2):
    """Full comparison of one image, decomposed the way north_star states the tolerances.

    pred_got / pred_ref: decoded predictions (rows, 5+nc) of the CUDA path / the oracle; dets_got: the CUDA path's detections.
      1. decode parity: max |obj, cls| difference over all rows (`conf_err`, tolerance 5e-3);
      2. NMS on IDENTICAL input: the oracle's non_max_suppression (genuine torchvision.ops.nms) run on the CUDA path's own
         predictions must reproduce its detections bit for bit (`nms_exact`) -- this also recovers their source rows;
      3. end to end: detections are matched BY SOURCE ROW (the same anchor cell on both sides, so nothing can pair up with
         a neighbour): `box_err` (max |xyxy| difference / box diagonal, tolerance 1e-2) and `det_conf_err` (5e-3) over the
         rows both sides keep; a row only one side keeps must be a threshold flip of upstream's own rules, proven from the
         OTHER side's numbers: its confidence there lies within conf_tol of conf_thres (`flips_conf`); or the box that
         suppresses it there overlaps it with an IoU within iou_tol of iou_thres (`flips_iou`); or that box outranks it by
         less than 2 conf_tol, so the pair swapped order (`flips_order`); or that box is itself one-sided (`flips_cascade`,
         the consequence of a flip counted elsewhere); or its best class differs between the sides while the row's two
         best class scores are within 2 conf_tol (`flips_class`: class-aware NMS then compares it with other boxes).
         Anything else is `unexplained`.
    """
    from . import nms as N
    pg = torch.as_tensor(np.asarray(pred_got)).float()
    pr = torch.as_tensor(np.asarray(pred_ref)).float()
    dg = torch.as_tensor(np.asarray(dets_got)).float()
    out = dict(conf_err=float((pg[:, 4:] - pr[:, 4:]).abs().max()), nms_exact=False, box_err=0.0, det_conf_err=0.0, common=0,
               box_over_tol=0, class_mismatch=0, box_err_worst=None,
               flips_conf=0, flips_iou=0, flips_order=0, flips_cascade=0, flips_class=0, unexplained=0, n_got=len(dg), n_ref=0)
    kw = dict(conf_thres=conf_thres, iou_thres=iou_thres, max_det=max_det, return_rows=True)
    (og,), (rows_g,) = N.non_max_suppression(pg[None], **kw)
    out["nms_exact"] = bool(og.shape == dg.shape and torch.equal(og, dg))
    (orf,), (rows_r,) = N.non_max_suppression(pr[None], **kw)
    out["n_ref"] = len(orf)
    side = {}
    for name, p, dets, rows in (("g", pg, og, rows_g), ("r", pr, orf, rows_r)):
        box, conf, cls, crow = N.candidates(p, conf_thres)
        side[name] = dict(kept={int(r): d for r, d in zip(rows.tolist(), dets)}, cand={int(r): i for i, r in enumerate(crow.tolist())},
                          box=box, conf=conf, cls=cls, p=p)
    both = sorted(set(side["g"]["kept"]) & set(side["r"]["kept"]))
    if both:
        a = torch.stack([side["g"]["kept"][r] for r in both])
        b = torch.stack([side["r"]["kept"][r] for r in both])
        diag = (b[:, 2:4] - b[:, 0:2]).norm(dim=1).clamp(min=1.0)
        rel = (a[:, :4] - b[:, :4]).abs().max(1).values / diag
        w = int(rel.argmax())
        out["box_err"] = float(rel[w])
        out["box_err_worst"] = dict(abs_px=float((a[w, :4] - b[w, :4]).abs().max()), diag_px=float(diag[w]), conf=float(b[w, 4]))
        out["box_over_tol"] = int((rel >= 1e-2).sum())
        out["det_conf_err"] = float((a[:, 4] - b[:, 4]).abs().max())
        out["class_mismatch"] = int((a[:, 5] != b[:, 5]).sum())
        for r in [both[j] for j in torch.nonzero(a[:, 5] != b[:, 5]).flatten().tolist()]:  # kept on both sides, other class
            sc = torch.sort(side["r"]["p"][r][5:] * side["r"]["p"][r][4], descending=True).values
            out["flips_class" if len(sc) > 1 and float(sc[0] - sc[1]) < 2 * conf_tol else "unexplained"] += 1
        out["common"] = len(both)
    lone = {"g": set(side["g"]["kept"]) - set(side["r"]["kept"]), "r": set(side["r"]["kept"]) - set(side["g"]["kept"])}
    for s, t in (("g", "r"), ("r", "g")):
        S, T = side[s], side[t]
        for r in lone[s]:
            d = S["kept"][r]
            if r not in T["cand"]:  # not even a candidate on the other side: its confidence there must sit at the threshold
                x = T["p"][r]
                c_t = float((x[5:] * x[4]).max()) if float(x[4]) > 0 else 0.0
                ok = abs(c_t - conf_thres) < conf_tol or abs(float(x[4]) - conf_thres) < conf_tol or abs(float(d[4]) - conf_thres) < conf_tol
                out["flips_conf" if ok else "unexplained"] += 1
                continue
            i = T["cand"][r]
            if float(d[5]) != float(T["cls"][i]):
                # the best class differs between the sides (class-aware NMS then sees another box): legitimate only when
                # the two best class scores of the row are within noise of each other
                sc = torch.sort(T["p"][r][5:] * T["p"][r][4], descending=True).values
                out["flips_class" if len(sc) > 1 and float(sc[0] - sc[1]) < 2 * conf_tol else "unexplained"] += 1
                continue
            # kept boxes of the other side that outrank or nearly outrank it, same class, overlapping
            kr = [k for k in T["kept"] if float(T["kept"][k][5]) == float(T["cls"][i])]
            why = None
            if kr:
                kb = torch.stack([T["kept"][k] for k in kr])
                v = box_iou(T["box"][i][None], kb[:, :4])[0]
                for j, k in enumerate(kr):
                    if v[j] <= iou_thres - iou_tol:
                        continue
                    if abs(float(v[j]) - iou_thres) < iou_tol:
                        why = why or "flips_iou"
                    elif abs(float(kb[j, 4]) - float(T["conf"][i])) < 2 * conf_tol:
                        why = why or "flips_order"
                    elif k in lone[t]:
                        why = why or "flips_cascade"
            if why is None and len(T["kept"]) >= max_det:
                why = "flips_order"  # cut by max_det on the other side
            out[why or "unexplained"] += 1
    return out
