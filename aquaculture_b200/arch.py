"""YOLOv5m graph description used by the host-side planner.

The reference runs `yolov5/detect.py` from the (absent) ultralytics/yolov5 submodule
[REF .SUBMODULES.json:8-16], invoked at [REF README.md:77].  The layer table below
restates upstream `models/yolov5m.yaml` + `models/yolo.py::parse_model`
[UPSTREAM-RECALLED, SURVEY.md App. A]: depth_multiple 0.67, width_multiple 0.75.

Only shapes and wiring live here; no arithmetic.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Sequence, Tuple

DEPTH_MULTIPLE = 0.67
WIDTH_MULTIPLE = 0.75

# default anchors in pixels (P3/8, P4/16, P5/32); a checkpoint may override them
DEFAULT_ANCHORS_PX: Tuple[Tuple[float, ...], ...] = (
    (10, 13, 16, 30, 33, 23),
    (30, 61, 62, 45, 59, 119),
    (116, 90, 156, 198, 373, 326),
)
STRIDES = (8, 16, 32)
BN_EPS = 1e-3  # upstream sets BatchNorm2d.eps = 1e-3 in Model._initialize / initialize_weights


def make_divisible(x: float, divisor: int = 8) -> int:
    return int(math.ceil(x / divisor) * divisor)


def _w(c: int) -> int:
    return make_divisible(c * WIDTH_MULTIPLE, 8)


def _d(n: int) -> int:
    return max(round(n * DEPTH_MULTIPLE), 1) if n > 1 else n


@dataclass
class Layer:
    idx: int
    kind: str                 # 'Conv' | 'C3' | 'SPPF' | 'Upsample' | 'Concat' | 'Detect'
    frm: Sequence[int]        # absolute source layer indices (-1 resolved)
    c_in: int
    c_out: int
    k: int = 1
    s: int = 1
    p: int = 0
    n: int = 1                # C3 repeats
    shortcut: bool = True


def yolov5m_layers(nc: int = 3) -> List[Layer]:
    """25 yaml layers of YOLOv5m (v6/v7 layout: 6x6 stem, SPPF)."""
    L: List[Layer] = []
    ch: List[int] = []

    def add(kind, frm, c_out, **kw):
        idx = len(L)
        frm_abs = [f if f >= 0 else idx + f for f in frm]
        if kind == "Concat":
            c_in = sum(ch[f] for f in frm_abs)
            c_out_ = c_in
        elif kind == "Upsample":
            c_in = ch[frm_abs[0]]
            c_out_ = c_in
        elif kind == "Detect":
            c_in = 0
            c_out_ = 0
        else:
            c_in = ch[frm_abs[0]] if idx > 0 else 3
            c_out_ = c_out
        L.append(Layer(idx, kind, frm_abs, c_in, c_out_, **kw))
        ch.append(c_out_)

    add("Conv", [-1], _w(64), k=6, s=2, p=2)          # 0  P1/2
    add("Conv", [-1], _w(128), k=3, s=2, p=1)         # 1  P2/4
    add("C3", [-1], _w(128), n=_d(3))                 # 2
    add("Conv", [-1], _w(256), k=3, s=2, p=1)         # 3  P3/8
    add("C3", [-1], _w(256), n=_d(6))                 # 4
    add("Conv", [-1], _w(512), k=3, s=2, p=1)         # 5  P4/16
    add("C3", [-1], _w(512), n=_d(9))                 # 6
    add("Conv", [-1], _w(1024), k=3, s=2, p=1)        # 7  P5/32
    add("C3", [-1], _w(1024), n=_d(3))                # 8
    add("SPPF", [-1], _w(1024), k=5)                  # 9
    add("Conv", [-1], _w(512), k=1, s=1, p=0)         # 10
    add("Upsample", [-1], 0)                          # 11
    add("Concat", [-1, 6], 0)                         # 12
    add("C3", [-1], _w(512), n=_d(3), shortcut=False)  # 13
    add("Conv", [-1], _w(256), k=1, s=1, p=0)         # 14
    add("Upsample", [-1], 0)                          # 15
    add("Concat", [-1, 4], 0)                         # 16
    add("C3", [-1], _w(256), n=_d(3), shortcut=False)  # 17 P3
    add("Conv", [-1], _w(256), k=3, s=2, p=1)         # 18
    add("Concat", [-1, 14], 0)                        # 19
    add("C3", [-1], _w(512), n=_d(3), shortcut=False)  # 20 P4
    add("Conv", [-1], _w(512), k=3, s=2, p=1)         # 21
    add("Concat", [-1, 10], 0)                        # 22
    add("C3", [-1], _w(1024), n=_d(3), shortcut=False)  # 23 P5
    add("Detect", [17, 20, 23], 0)                    # 24
    return L


def detect_rows(h: int, w: int) -> int:
    """Rows of the Detect output for an (h, w) network input: 3 anchors per cell per level."""
    return sum(3 * (h // s) * (w // s) for s in STRIDES)


def conv_flops_per_tile(h: int = 640, w: int = 640, nc: int = 3) -> float:
    """Algorithmic conv FLOPs (2*MAC) of the 82 convs for one (h, w) tile."""
    total = 0.0
    for spec in iter_convs(nc):
        ho, wo = h // spec.out_stride, w // spec.out_stride
        total += 2.0 * ho * wo * spec.c_out * spec.c_in * spec.k * spec.k
    return total


@dataclass
class ConvSpec:
    name: str        # upstream state_dict prefix, e.g. 'model.2.m.0.cv1'
    c_in: int
    c_out: int
    k: int
    s: int
    p: int
    out_stride: int  # network stride of the OUTPUT map
    act: bool = True  # Conv+BN+SiLU (True) or plain biased Conv2d (Detect head)


def iter_convs(nc: int = 3) -> List[ConvSpec]:
    """Every conv in execution order with its upstream parameter prefix (82 for YOLOv5m)."""
    out: List[ConvSpec] = []
    stride_of = {}
    cur = 1
    for l in yolov5m_layers(nc):
        pre = f"model.{l.idx}"
        if l.kind == "Conv":
            src = stride_of.get(l.frm[0], 1) if l.idx > 0 else 1
            cur = src * l.s
            out.append(ConvSpec(pre, l.c_in, l.c_out, l.k, l.s, l.p, cur))
        elif l.kind == "C3":
            cur = stride_of[l.frm[0]]
            c_ = l.c_out // 2
            out.append(ConvSpec(pre + ".cv1", l.c_in, c_, 1, 1, 0, cur))
            out.append(ConvSpec(pre + ".cv2", l.c_in, c_, 1, 1, 0, cur))
            for j in range(l.n):
                out.append(ConvSpec(f"{pre}.m.{j}.cv1", c_, c_, 1, 1, 0, cur))
                out.append(ConvSpec(f"{pre}.m.{j}.cv2", c_, c_, 3, 1, 1, cur))
            out.append(ConvSpec(pre + ".cv3", 2 * c_, l.c_out, 1, 1, 0, cur))
        elif l.kind == "SPPF":
            cur = stride_of[l.frm[0]]
            c_ = l.c_in // 2
            out.append(ConvSpec(pre + ".cv1", l.c_in, c_, 1, 1, 0, cur))
            out.append(ConvSpec(pre + ".cv2", 4 * c_, l.c_out, 1, 1, 0, cur))
        elif l.kind == "Upsample":
            cur = stride_of[l.frm[0]] // 2
        elif l.kind == "Concat":
            cur = stride_of[l.frm[0]]
        elif l.kind == "Detect":
            layers = yolov5m_layers(nc)
            for j, f in enumerate(l.frm):
                out.append(ConvSpec(f"{pre}.m.{j}", layers[f].c_out, 3 * (nc + 5), 1, 1, 0,
                                    stride_of[f], act=False))
        stride_of[l.idx] = cur
    return out
