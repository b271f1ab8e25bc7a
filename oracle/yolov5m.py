"""ORACLE (test infrastructure, never on the product path): fp32 CPU restatement of YOLOv5m.

PARITY UNPINNED for the model graph: the reference's implementation is the absent
ultralytics/yolov5 submodule @1db95338 [REF .SUBMODULES.json:8-16] and its weights are a missing
blob [REF .MISSING_LARGE_BLOBS:3]; no golden model outputs exist.  This restates, from SURVEY.md
§8a/App. A [UPSTREAM-RECALLED]:
  models/common.py::{autopad, Conv, Bottleneck, C3, SPPF, Concat}
  models/yolo.py::{Detect.forward (eval branch), Detect._make_grid, DetectionModel._forward_once}
  utils/torch_utils.py::fuse_conv_and_bn (via aquaculture_b200.weights.fold_bn formulas, re-derived here)
What pins it: torch.nn.functional conv2d/max_pool2d/interpolate (genuine third-party kernels) and
the architecture self-check 21.17 M params / 48.87 GFLOPs at nc=80 (tests/test_oracle.py).
"""
from __future__ import annotations

import torch
import torch.nn as nn

YAML = dict(
    depth_multiple=0.67, width_multiple=0.75,
    anchors=[[10, 13, 16, 30, 33, 23], [30, 61, 62, 45, 59, 119], [116, 90, 156, 198, 373, 326]],
    backbone=[[-1, 1, "Conv", [64, 6, 2, 2]], [-1, 1, "Conv", [128, 3, 2]], [-1, 3, "C3", [128]],
              [-1, 1, "Conv", [256, 3, 2]], [-1, 6, "C3", [256]], [-1, 1, "Conv", [512, 3, 2]],
              [-1, 9, "C3", [512]], [-1, 1, "Conv", [1024, 3, 2]], [-1, 3, "C3", [1024]],
              [-1, 1, "SPPF", [1024, 5]]],
    head=[[-1, 1, "Conv", [512, 1, 1]], [-1, 1, "Upsample", [None, 2, "nearest"]],
          [[-1, 6], 1, "Concat", [1]], [-1, 3, "C3", [512, False]], [-1, 1, "Conv", [256, 1, 1]],
          [-1, 1, "Upsample", [None, 2, "nearest"]], [[-1, 4], 1, "Concat", [1]],
          [-1, 3, "C3", [256, False]], [-1, 1, "Conv", [256, 3, 2]], [[-1, 14], 1, "Concat", [1]],
          [-1, 3, "C3", [512, False]], [-1, 1, "Conv", [512, 3, 2]], [[-1, 10], 1, "Concat", [1]],
          [-1, 3, "C3", [1024, False]], [[17, 20, 23], 1, "Detect", ["nc", "anchors"]]],
)


def autopad(k, p=None):
    return k // 2 if p is None else p


class Conv(nn.Module):
    def __init__(self, c1, c2, k=1, s=1, p=None):
        super().__init__()
        self.conv = nn.Conv2d(c1, c2, k, s, autopad(k, p), bias=False)
        self.bn = nn.BatchNorm2d(c2, eps=1e-3, momentum=0.03)
        self.act = nn.SiLU()

    def forward(self, x):
        return self.act(self.bn(self.conv(x)))


class Bottleneck(nn.Module):
    def __init__(self, c1, c2, shortcut=True, e=0.5):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c_, c2, 3, 1)
        self.add = shortcut and c1 == c2

    def forward(self, x):
        return x + self.cv2(self.cv1(x)) if self.add else self.cv2(self.cv1(x))


class C3(nn.Module):
    def __init__(self, c1, c2, n=1, shortcut=True, e=0.5):
        super().__init__()
        c_ = int(c2 * e)
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c1, c_, 1, 1)
        self.cv3 = Conv(2 * c_, c2, 1)
        self.m = nn.Sequential(*(Bottleneck(c_, c_, shortcut, e=1.0) for _ in range(n)))

    def forward(self, x):
        return self.cv3(torch.cat((self.m(self.cv1(x)), self.cv2(x)), 1))


class SPPF(nn.Module):
    def __init__(self, c1, c2, k=5):
        super().__init__()
        c_ = c1 // 2
        self.cv1 = Conv(c1, c_, 1, 1)
        self.cv2 = Conv(c_ * 4, c2, 1, 1)
        self.m = nn.MaxPool2d(kernel_size=k, stride=1, padding=k // 2)

    def forward(self, x):
        x = self.cv1(x)
        y1 = self.m(x)
        y2 = self.m(y1)
        return self.cv2(torch.cat((x, y1, y2, self.m(y2)), 1))


class Concat(nn.Module):
    def forward(self, xs):
        return torch.cat(xs, 1)


class Detect(nn.Module):
    def __init__(self, nc, anchors, ch):
        super().__init__()
        self.nc, self.no, self.nl, self.na = nc, nc + 5, len(anchors), len(anchors[0]) // 2
        self.register_buffer("anchors", torch.tensor(anchors).float().view(self.nl, -1, 2))
        self.m = nn.ModuleList(nn.Conv2d(x, self.no * self.na, 1) for x in ch)
        self.stride = torch.tensor([8.0, 16.0, 32.0])

    def forward(self, xs):
        z = []
        for i, x in enumerate(xs):
            x = self.m[i](x)
            bs, _, ny, nx = x.shape
            x = x.view(bs, self.na, self.no, ny, nx).permute(0, 1, 3, 4, 2).contiguous()
            yv, xv = torch.meshgrid(torch.arange(ny, dtype=x.dtype), torch.arange(nx, dtype=x.dtype), indexing="ij")
            grid = torch.stack((xv, yv), 2).expand(1, self.na, ny, nx, 2) - 0.5
            anchor_grid = (self.anchors[i] * self.stride[i]).view(1, self.na, 1, 1, 2).expand(1, self.na, ny, nx, 2)
            xy, wh, conf = x.sigmoid().split((2, 2, self.nc + 1), 4)
            xy = (xy * 2 + grid) * self.stride[i]
            wh = (wh * 2) ** 2 * anchor_grid
            z.append(torch.cat((xy, wh, conf), 4).view(bs, self.na * nx * ny, self.no))
        return torch.cat(z, 1)


def _make_divisible(x, d=8):
    import math
    return math.ceil(x / d) * d


class YOLOv5m(nn.Module):
    """`DetectionModel` restated: parse the yaml into 25 modules and run `_forward_once`."""

    def __init__(self, nc=3):
        super().__init__()
        gd, gw = YAML["depth_multiple"], YAML["width_multiple"]
        ch, layers, self.frm, self.save = [3], [], [], set()
        for i, (f, n, m, args) in enumerate(YAML["backbone"] + YAML["head"]):
            n = max(round(n * gd), 1) if n > 1 else n
            if m in ("Conv", "C3", "SPPF"):
                c1, c2 = ch[f], _make_divisible(args[0] * gw, 8)
                if m == "Conv":
                    mod = Conv(c1, c2, *args[1:])
                elif m == "C3":
                    mod = C3(c1, c2, n, *args[1:])
                else:
                    mod = SPPF(c1, c2, *args[1:])
            elif m == "Upsample":
                mod, c2 = nn.Upsample(None, 2, "nearest"), ch[f]
            elif m == "Concat":
                mod, c2 = Concat(), sum(ch[x] for x in f)
            else:
                mod, c2 = Detect(nc, YAML["anchors"], [ch[x] for x in f]), 0
            layers.append(mod)
            self.frm.append(f)
            self.save.update(x % i for x in ([f] if isinstance(f, int) else f) if x != -1)
            if i == 0:
                ch = []
            ch.append(c2)
        self.model = nn.Sequential(*layers)
        self.nc = nc

    def forward(self, x):
        y = []
        for m, f in zip(self.model, self.frm):
            if f != -1:
                x = y[f] if isinstance(f, int) else [x if j == -1 else y[j] for j in f]
            x = m(x)
            y.append(x)
        return x

    def fuse(self):
        """fuse_conv_and_bn on every Conv: conv gets a bias, bn is dropped."""
        for m in self.modules():
            if isinstance(m, Conv) and hasattr(m, "bn"):
                w = m.conv.weight.detach()
                scale = m.bn.weight.detach() / torch.sqrt(m.bn.running_var + m.bn.eps)
                fused = nn.Conv2d(m.conv.in_channels, m.conv.out_channels, m.conv.kernel_size, m.conv.stride,
                                  m.conv.padding, bias=True)
                fused.weight.data.copy_((torch.diag(scale) @ w.view(w.shape[0], -1)).view(w.shape))
                fused.bias.data.copy_(m.bn.bias.detach() - m.bn.weight.detach() * m.bn.running_mean / torch.sqrt(m.bn.running_var + m.bn.eps))
                m.conv = fused
                del m.bn
                m.forward = (lambda mm: (lambda x: mm.act(mm.conv(x))))(m)
        return self


def build(state_dict, nc=3, fuse=True):
    """Oracle model in eval mode with the given upstream-named state dict loaded."""
    model = YOLOv5m(nc)
    sd = {k: v for k, v in state_dict.items()}
    missing, unexpected = model.load_state_dict(sd, strict=False)
    missing = [k for k in missing if "num_batches_tracked" not in k]
    assert not missing and not unexpected, (missing, unexpected)
    model.eval()
    return model.fuse() if fuse else model
