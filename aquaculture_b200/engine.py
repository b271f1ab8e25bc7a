"""Host-side mirror of the reference's model wiring for the tile path.

`TileDetector` plays the role of upstream `DetectMultiBackend` + `DetectionModel._forward_once` +
`non_max_suppression` as `yolov5/detect.py` drives them [REF README.md:77; UPSTREAM-RECALLED SURVEY.md
§3.1-3.2], but it only *plans* the graph: every arithmetic step runs in libaqua_b200.so (hand-written
sm_100a CUDA behind the C ABI in include/aqua_b200.h).  torch is used for weight tensors and pinned
host memory only.  There is no CPU or torch execution path; a missing library or GPU raises.

Graph rewrites done here (all exact):
  * BN folded into conv weights/bias in fp32, then cast to fp16 (`weights.fold_bn`);
  * 6x6/s2 stem == 3x3/s1 conv over the 2x2 space-to-depth input (12 -> 16 channels), which the
    preprocess kernel writes directly;
  * C3.cv1 and C3.cv2 read the same tensor -> one GEMM with N = 2c_;
  * torch.cat / Concat never copy: producers write into channel slices of the consumer's buffer;
  * optionally (`stage_plan`, AQUA_STAGES) the high-resolution layers run depth-first over sub-batches, their
    activations produced and consumed inside the 126 MB L2.  Same kernels, same values, another launch order.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib, arch, weights
from ._lib import AQ_ACT_NONE, AQ_ACT_SILU, AQ_ACT_SILU_FAST, AquaError, ConvDesc, DetectDesc

MAX_WH = 7680.0   # upstream non_max_suppression: class offset (pixels)
MAX_NMS = 30000   # upstream non_max_suppression: boxes into torchvision.ops.nms


def pack_krsc(w_oihw: torch.Tensor) -> np.ndarray:
    """OIHW fp32 -> [O][kh][kw][I] fp16 bits (uint16), contiguous."""
    return np.ascontiguousarray(w_oihw.permute(0, 2, 3, 1).contiguous().half().numpy().view(np.uint16))


def pack_stem_s2d(w_oihw: torch.Tensor) -> torch.Tensor:
    """6x6/s2/p2 stem weights (O,3,6,6) -> equivalent 3x3/s1/p1 weights (O,16,3,3) over the space-to-depth
    input whose channel is (py*2+px)*3 + rgb: W2[o, (py*2+px)*3+c, a, b] = W[o, c, 2a+py, 2b+px]."""
    o = w_oihw.shape[0]
    w2 = torch.zeros(o, 16, 3, 3, dtype=w_oihw.dtype)
    for py in range(2):
        for px in range(2):
            for c in range(3):
                w2[:, (py * 2 + px) * 3 + c] = w_oihw[:, c, py::2, px::2]
    return w2


class Context:
    """Owns one aq_ctx."""

    def __init__(self, device: int = 0):
        self.lib = _lib.load()
        h = C.c_void_p()
        rc = self.lib.aq_create(device, C.byref(h))
        if rc < 0:
            raise AquaError(f"aq_create(device={device}) failed ({rc}): {self.lib.aq_last_error(None).decode()}")
        self.h = h
        self.device = device
        self._keep = []  # host arrays referenced by descriptors until plan_end
        # the shipped tiling table: identical tilings (hence bit-identical label text) in every process and rank
        self.tuned_entries = 0
        if os.path.exists(_lib.TUNE_TABLE) and os.environ.get("AQUA_TUNE_TABLE", "") != "0":
            self.tuned_entries = self.call("aq_tune_load", _lib.TUNE_TABLE.encode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.aq_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def call(self, name: str, *args) -> int:
        return _lib.check(self.lib, self.h, getattr(self.lib, name)(self.h, *args), name)

    # ---- plan helpers
    def plan_begin(self, batch, net_h, net_w, nc):
        self.call("aq_plan_begin", batch, net_h, net_w, nc)

    def buffer(self, h, w, c) -> int:
        return self.call("aq_plan_buffer", h, w, c)

    def stage_begin(self, sub_batch: int, lanes: int = 1):
        """Ops and buffers planned until `stage_end` run depth-first over sub-batches (aq_plan_stage_begin)."""
        self.call("aq_plan_stage_begin", int(sub_batch), int(lanes))

    def stage_end(self):
        self.call("aq_plan_stage_end")

    def conv(self, src, dst, w_oihw: torch.Tensor, bias: torch.Tensor, k, stride, act=True, res=None,
             tile_n=0, mblk=0, mode=0, cluster=0, up=None):
        """src/dst/res are (buffer id, channel offset); w_oihw fp32 (O,I,k,k).  act: False/0 none, True/1 SiLU
        (ex2+rcp), 2 SiLU via tanh.approx; res must equal dst (in-place Bottleneck add).  up = (buffer id, channel
        offset, channels) of a half-resolution tensor whose nearest-2x upsampling forms the FIRST `channels` input
        channels (upstream Upsample + Concat folded into the conv's TMA loads); src then supplies the rest."""
        wk = pack_krsc(w_oihw)
        b = np.ascontiguousarray(bias.float().numpy())
        self._keep += [wk, b]
        d = ConvDesc(src[0], src[1], w_oihw.shape[1], dst[0], dst[1], w_oihw.shape[0], k, stride,
                     int(act), res[0] if res else -1, res[1] if res else 0,
                     wk.ctypes.data, b.ctypes.data, tile_n, mblk, mode, cluster,
                     up[0] if up else -1, up[1] if up else 0, up[2] if up else 0)
        self.call("aq_plan_conv", C.byref(d))

    def conv_chain(self, src, dst, w_oihw, bias, k, stride, act, dst2, w2_oihw, bias2, act2, store_first=True, up=None,
                   mblk=0, mode=0):  # noqa: D401
        """`conv(src -> dst)` followed, inside the SAME kernel, by a 1x1 conv over the first `w2.shape[1]` channels of its
        output into dst2 (aq_plan_conv_chain).  Raises AquaError (AQ_ERR_UNSUPPORTED) when no tiling can host the pair."""
        wk, wk2 = pack_krsc(w_oihw), pack_krsc(w2_oihw)
        b, b2 = np.ascontiguousarray(bias.float().numpy()), np.ascontiguousarray(bias2.float().numpy())
        self._keep += [wk, wk2, b, b2]
        d1 = ConvDesc(src[0], src[1], w_oihw.shape[1], dst[0], dst[1], w_oihw.shape[0], k, stride, int(act), -1, 0,
                      wk.ctypes.data, b.ctypes.data, 0, mblk, mode, 0, up[0] if up else -1, up[1] if up else 0, up[2] if up else 0)
        d2 = ConvDesc(dst[0], dst[1], w2_oihw.shape[1], dst2[0], dst2[1], w2_oihw.shape[0], 1, 1, int(act2), -1, 0,
                      wk2.ctypes.data, b2.ctypes.data, 0, 0, 0, 0, -1, 0, 0)
        self.call("aq_plan_conv_chain", C.byref(d1), C.byref(d2), int(bool(store_first)))

    def detect_level(self, src, w_oi: torch.Tensor, bias: torch.Tensor, nc, level_row_off, stride, anchors_px):
        wk = np.ascontiguousarray(w_oi.reshape(w_oi.shape[0], -1).half().numpy().view(np.uint16))
        b = np.ascontiguousarray(bias.float().numpy())
        self._keep += [wk, b]
        d = DetectDesc(src[0], src[1], wk.shape[1], nc, level_row_off, float(stride),
                       (C.c_float * 6)(*[float(a) for a in anchors_px]), wk.ctypes.data, b.ctypes.data)
        self.call("aq_plan_detect", C.byref(d))

    def plan_end(self):
        self.call("aq_plan_end")
        self._keep = []

    def buffer_info(self, buf) -> Tuple[int, Tuple[int, int, int, int]]:
        p = C.c_void_p()
        n, h, w, c = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self.call("aq_buffer_info", buf, C.byref(p), C.byref(n), C.byref(h), C.byref(w), C.byref(c))
        return p.value, (n.value, h.value, w.value, c.value)

    def upload(self, buf: int, arr: np.ndarray):
        """NHWC fp16 numpy -> planned buffer."""
        ptr, shape = self.buffer_info(buf)
        assert tuple(arr.shape) == shape and arr.dtype == np.float16, (arr.shape, shape, arr.dtype)
        arr = np.ascontiguousarray(arr)
        self.call("aq_memcpy", C.c_void_p(ptr), arr.ctypes.data, arr.nbytes, 0, None)
        self.call("aq_sync", None)

    def download(self, buf: int) -> np.ndarray:
        ptr, shape = self.buffer_info(buf)
        out = np.empty(shape, np.float16)
        self.call("aq_memcpy", out.ctypes.data, C.c_void_p(ptr), out.nbytes, 1, None)
        self.call("aq_sync", None)
        return out

    def forward(self):
        self.call("aq_forward", None)

    def sync(self):
        self.call("aq_sync", None)

    def pred(self) -> np.ndarray:
        p = C.c_void_p()
        rows, no = C.c_int(), C.c_int()
        self.call("aq_pred_info", C.byref(p), C.byref(rows), C.byref(no))
        _, (n, _, _, _) = self.buffer_info(0)
        out = np.empty((n, rows.value, no.value), np.float32)
        self.call("aq_memcpy", out.ctypes.data, p, out.nbytes, 1, None)
        self.call("aq_sync", None)
        return out


def nms_device(ctx: Context, pred: torch.Tensor, conf_thres=0.25, iou_thres=0.45, max_det=1000, agnostic=False,
               classes: Optional[Sequence[int]] = None, max_wh=MAX_WH, max_nms=MAX_NMS):
    """Standalone batched NMS on a CUDA fp32 tensor (B, rows, no) -> (dets (B,max_det,6), counts (B), keep_rows)."""
    assert pred.is_cuda and pred.dtype == torch.float32 and pred.is_contiguous()
    b, rows, no = pred.shape
    dets = torch.zeros(b, max_det, 6, device=pred.device)
    counts = torch.zeros(b, dtype=torch.int32, device=pred.device)
    keep = torch.full((b, max_det), -1, dtype=torch.int32, device=pred.device)
    cls_arr = (C.c_int * len(classes))(*classes) if classes else None
    torch.cuda.synchronize()
    ctx.call("aq_nms", C.c_void_p(pred.data_ptr()), b, rows, no, float(conf_thres), float(iou_thres), max_det,
             int(agnostic), cls_arr, len(classes) if classes else 0, float(max_wh), max_nms,
             C.c_void_p(dets.data_ptr()), C.c_void_p(counts.data_ptr()), C.c_void_p(keep.data_ptr()), None)
    ctx.sync()
    return dets, counts, keep


@dataclass
class Detections:
    dets: np.ndarray    # (n, max_det, 6) fp32: x1,y1,x2,y2,conf,cls in letterboxed pixels, descending conf
    counts: np.ndarray  # (n,) int32


# Depth-first stages of the conv stack, "name=<sub-batch images at 640x640>x<lanes>".  a: stem .. layer 3 (320^2 / 160^2
# maps), b: C3 of layer 4 (80^2), c: C3 of layer 17 + its Detect level + layer 18 (80^2 head).  The sub-batch is scaled
# by 640*640 / (net_h * net_w) for other input sizes (at least one image).  AQUA_STAGES="a=8x1,b=16x2" turns stages on;
# "0" / unset plans every layer over the whole batch (upstream's order).
# OFF by default -- measured on a B200 (profiles/r2_stage_sweep_v1.txt): keeping the 320^2 / 160^2 activations in L2 does
# take ~40 % off those layers, but each extra launch costs 5-8 us of pipeline drain + fill (one 672-thread CTA per SM,
# nothing to overlap it with), so every split tried (sub-batches 2 .. 32, 1 .. 3 lanes) ends slower than the 75-launch
# plan: 4.47 ms unstaged vs 4.82 .. 5.86 ms staged per 64-tile batch.  The mechanism is exact (same kernels, same values;
# tests/test_gpu_kernels.py::test_staged_plan_*) and stays available for larger inputs.
DEFAULT_STAGES = {}
# Default: both kinds of chain (measured, interleaved A/B of the batch-64 conv stack, profiles/r2_ab_chain_serpentine.txt:
# 4.615 ms none, 4.580 "c3", 4.569 "feed", 4.550 "feed,c3").
DEFAULT_CHAIN = "feed,c3"


def stage_plan(net_h: int, net_w: int, batch: int, spec: Optional[str] = None) -> Dict[str, Tuple[int, int]]:
    """{stage: (sub-batch images, lanes)} for this input size; stages that would not split the batch are left out."""
    spec = os.environ.get("AQUA_STAGES", "") if spec is None else spec
    if spec.strip() == "0":
        return {}
    plan = dict(DEFAULT_STAGES)
    for item in filter(None, (t.strip() for t in spec.split(","))):
        name, val = item.split("=")
        sub, _, lanes = val.partition("x")
        plan[name.strip()] = (int(sub), int(lanes or 1))
    out = {}
    for name, (sub, lanes) in plan.items():
        if sub <= 0:
            continue
        sub = max(1, sub * 640 * 640 // (net_h * net_w))
        while batch % sub:  # a whole number of sub-batches
            sub -= 1
        if sub < batch:
            out[name] = (sub, max(1, min(lanes, 4, batch // sub)))
    return out


class TileDetector:
    """YOLOv5m over batches of equally-sized tiles on one B200."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], nc: int, batch: int = 64, imgsz=(640, 640),
                 device: int = 0, conf_thres=0.25, iou_thres=0.45, max_det=1000, agnostic=False,
                 classes: Optional[Sequence[int]] = None, ctx=None):
        self.nc, self.batch = nc, batch
        self.net_h, self.net_w = int(imgsz[0]), int(imgsz[1])
        self.conf_thres, self.iou_thres, self.max_det = float(conf_thres), float(iou_thres), int(max_det)
        self.agnostic, self.classes = bool(agnostic), list(classes) if classes else None
        # SiLU evaluation in the conv epilogues: "fast" = tanh.approx form (<= 1.1e-5 absolute error, measured),
        # "exact" = ex2 + rcp.  Both are device code; this only selects the epilogue variant.
        self.act = AQ_ACT_SILU if os.environ.get("AQUA_SILU", "fast") == "exact" else AQ_ACT_SILU_FAST
        self.ctx = ctx if ctx is not None else Context(device)   # `ctx` lets tests inject the plan emulator
        self._build(state_dict)
        self.stage_ms = [0.0, 0.0, 0.0]  # device time of pre-process / inference / NMS summed over the waited batches (not the first)
        self.first_stage_ms, self.first_stage_tiles = [0.0, 0.0, 0.0], 0
        self.stage_tiles = 0
        self._dets = self._counts = None  # pinned host staging, double-buffered, allocated on first use
        self._slot = 0
        self._inflight = [None, None]
        self._pin, self._pin_i = {}, {}
        self._jpeg_out = {}

    # ------------------------------------------------------------------ plan
    def _build(self, sd):
        ctx, nc = self.ctx, self.nc
        H, W, B = self.net_h, self.net_w, self.batch
        fw = weights.fold_bn(sd, nc)
        ctx.plan_begin(B, H, W, nc)
        buf = ctx.buffer
        stages = stage_plan(H, W, B)
        self.stages = stages
        # 1x1 convs chained behind their producers (EPI_CHAIN): "c3", "feed", "feed,c3" ("1" = both), "0" = none
        chain_on = os.environ.get("AQUA_CHAIN", DEFAULT_CHAIN)
        chain_on = {"0": "", "1": "feed,c3"}.get(chain_on, chain_on)
        self.chained = []
        self.unwritten = set()  # buffers whose producer was chained with store_first = 0: they are never written

        class stage:  # `with stage("a"):` -- ops and buffers planned inside form one depth-first stage
            def __init__(self, name):
                self.cfg = stages.get(name)

            def __enter__(self):
                if self.cfg:
                    ctx.stage_begin(*self.cfg)

            def __exit__(self, *exc):
                if self.cfg and exc[0] is None:
                    ctx.stage_end()

        b_in = buf(H // 2, W // 2, 16)
        x3 = buf(H // 8, W // 8, 192)
        cat16 = buf(H // 8, W // 8, 384)    # [up(L14) | L4]
        x5 = buf(H // 16, W // 16, 384)
        cat12 = buf(H // 16, W // 16, 768)  # [up(L10) | L6]
        x7 = buf(H // 32, W // 32, 768)
        x8 = buf(H // 32, W // 32, 768)
        sppf = buf(H // 32, W // 32, 1536)
        x9 = buf(H // 32, W // 32, 768)
        cat22 = buf(H // 32, W // 32, 768)  # [L21 | L10]
        x13 = buf(H // 16, W // 16, 384)
        cat19 = buf(H // 16, W // 16, 384)  # [L18 | L14]
        x20 = buf(H // 16, W // 16, 384)
        x23 = buf(H // 32, W // 32, 768)
        assert b_in == 0

        def conv(name, src, dst, k, s, res=None):
            w, b = fw[name]
            ctx.conv(src, dst, w, b, k, s, self.act, res)

        def c3(idx, src, c1, dst, c2, n, shortcut, hw, up=None, feeder=None):
            """feeder = (name, src, k, s): the Conv layer whose ONLY consumer is this C3 (its output is `src`); with chaining on it
            runs in one kernel with cv1 || cv2 and its output never goes to memory."""
            c_ = c2 // 2
            cat = buf(hw[0], hw[1], 2 * c_)
            tmp = buf(hw[0], hw[1], c_)
            pre = f"model.{idx}"
            w1, b1 = fw[pre + ".cv1"]
            w2, b2 = fw[pre + ".cv2"]
            w12, b12 = torch.cat((w1, w2), 0), torch.cat((b1, b2), 0)
            # cv1 || cv2 is one GEMM.  Chaining (aq_plan_conv_chain, EPI_CHAIN): either the layer in front of the C3 runs in the
            # same kernel as cv1 || cv2 ("feed": its output is consumed out of shared memory and never stored), or the first
            # Bottleneck's 1x1 runs behind cv1 || cv2 ("c3").  Each saves a launch and a trip of the tensor through HBM / L2.
            done12 = first_chained = False
            if feeder is not None:
                fname, fsrc, fk, fs = feeder
                wf, bf = fw[fname]
                if "feed" in chain_on and 2 * c_ <= 256 and wf.shape[0] <= 256:
                    try:
                        ctx.conv_chain(fsrc, src, wf, bf, fk, fs, self.act, (cat, 0), w12, b12, self.act, store_first=False)
                        done12 = True
                        self.chained.append(f"{fname} -> {pre}.cv1||cv2")
                        self.unwritten.add(src[0])
                    except AquaError:
                        pass
                if not done12:
                    ctx.conv(fsrc, src, wf, bf, fk, fs, self.act)
            if not done12 and "c3" in chain_on and 2 * c_ <= 256:
                wm, bm = fw[f"{pre}.m.0.cv1"]
                try:
                    ctx.conv_chain(src, (cat, 0), w12, b12, 1, 1, self.act, (tmp, 0), wm, bm, self.act, store_first=True, up=up)
                    done12 = first_chained = True
                    self.chained.append(f"{pre}.cv1||cv2 -> m.0.cv1")
                except AquaError:
                    pass  # no tiling leaves room for the chained weights / accumulators: two launches
            if not done12:
                ctx.conv(src, (cat, 0), w12, b12, 1, 1, self.act, up=up)
            for j in range(n):
                if not (j == 0 and first_chained):
                    conv(f"{pre}.m.{j}.cv1", (cat, 0), (tmp, 0), 1, 1)
                conv(f"{pre}.m.{j}.cv2", (tmp, 0), (cat, 0), 3, 1, res=(cat, 0) if shortcut else None)
            conv(pre + ".cv3", (cat, 0), dst, 1, 1)

        anc = weights.anchors_px(sd)
        level_rows = [3 * (H // s) * (W // s) for s in (8, 16, 32)]
        level_off = [0, level_rows[0], level_rows[0] + level_rows[1]]

        def detect(lvl, src):
            w, b = fw[f"model.24.m.{lvl}"]
            ctx.detect_level((src, 0), w, b, nc, level_off[lvl], (8, 16, 32)[lvl], anc[lvl].reshape(-1))

        L = {l.idx: l for l in arch.yolov5m_layers(nc)}
        # backbone
        with stage("a"):  # 320^2 and 160^2 maps: the largest activations of the network
            x0 = buf(H // 2, W // 2, 48)
            x1 = buf(H // 4, W // 4, 96)
            x2 = buf(H // 4, W // 4, 96)
            w0, b0 = fw["model.0"]
            ctx.conv((b_in, 0), (x0, 0), pack_stem_s2d(w0), b0, 3, 1, self.act)
            c3(2, (x1, 0), 96, (x2, 0), 96, L[2].n, True, (H // 4, W // 4), feeder=("model.1", (x0, 0), 3, 2))
            if stages.get("a") or stages.get("b"):  # layer 3 and the C3 behind it sit in different stages: not chained
                conv("model.3", (x2, 0), (x3, 0), 3, 2)
        with stage("b"):
            c3(4, (x3, 0), 192, (cat16, 192), 192, L[4].n, True, (H // 8, W // 8),
               feeder=None if (stages.get("a") or stages.get("b")) else ("model.3", (x2, 0), 3, 2))
        conv("model.5", (cat16, 192), (x5, 0), 3, 2)
        c3(6, (x5, 0), 384, (cat12, 384), 384, L[6].n, True, (H // 16, W // 16))
        conv("model.7", (cat12, 384), (x7, 0), 3, 2)
        c3(8, (x7, 0), 768, (x8, 0), 768, L[8].n, True, (H // 32, W // 32))
        conv("model.9.cv1", (x8, 0), (sppf, 0), 1, 1)
        ctx.call("aq_plan_sppf_pool", sppf, 0, 384)
        conv("model.9.cv2", (sppf, 0), (x9, 0), 1, 1)
        # head
        conv("model.10", (x9, 0), (cat22, 384), 1, 1)
        # Upsample + Concat (layers 11-12, 15-16): folded into the consuming C3's first conv, whose TMA loads read the
        # half-resolution tensor through a stride-0 replicating tensor map (AQUA_FOLD_UPSAMPLE=0: separate kernel)
        fold = os.environ.get("AQUA_FOLD_UPSAMPLE", "1") != "0"
        if fold:
            c3(13, (cat12, 384), 768, (x13, 0), 384, L[13].n, False, (H // 16, W // 16), up=(cat22, 384, 384))
        else:
            ctx.call("aq_plan_upsample2x", cat22, 384, 384, cat12, 0)
            c3(13, (cat12, 0), 768, (x13, 0), 384, L[13].n, False, (H // 16, W // 16))
        conv("model.14", (x13, 0), (cat19, 192), 1, 1)
        if not fold:
            ctx.call("aq_plan_upsample2x", cat19, 192, 192, cat16, 0)
        with stage("c"):  # the 80^2 head: C3 of layer 17, its Detect level and the stride-2 conv that consumes it
            x17 = buf(H // 8, W // 8, 192)
            if fold:
                c3(17, (cat16, 192), 384, (x17, 0), 192, L[17].n, False, (H // 8, W // 8), up=(cat19, 192, 192))
            else:
                c3(17, (cat16, 0), 384, (x17, 0), 192, L[17].n, False, (H // 8, W // 8))
            detect(0, x17)
            conv("model.18", (x17, 0), (cat19, 0), 3, 2)
        c3(20, (cat19, 0), 384, (x20, 0), 384, L[20].n, False, (H // 16, W // 16))
        detect(1, x20)  # each Detect level right behind the layer that feeds it (measured: no different from running it last)
        conv("model.21", (x20, 0), (cat22, 0), 3, 2)
        c3(23, (cat22, 0), 768, (x23, 0), 768, L[23].n, False, (H // 32, W // 32))
        detect(2, x23)
        self.rows = sum(level_rows)
        ctx.plan_end()
        self.n_launches = ctx.call("aq_plan_num_launches")
        self.buffers = dict(x0=x0, x1=x1, x2=x2, x3=x3, x17=x17, x20=x20, x23=x23, x9=x9, x8=x8)

    # ------------------------------------------------------------------ run
    def detect_submit(self, tiles, bgr=True, src_is_device=False, stream=None):
        """Enqueue one batch.  tiles: uint8 (n, h, w, 3) torch tensor (pinned host, or cuda).  Returns a ticket;
        `detect_wait(ticket)` blocks until that batch's detections are in pinned host memory.  Two batches may be
        in flight: the H2D copy of the second overlaps the kernels of the first (the tiles tensor of a batch must
        stay alive and unmodified until its ticket has been waited for)."""
        n, h, w, _ = tiles.shape
        assert n <= self.batch and tiles.dtype == torch.uint8 and tiles.is_contiguous()
        if self._dets is None:
            self._dets = [torch.empty(self.batch, self.max_det, 6, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._counts = [torch.empty(self.batch, dtype=torch.int32).pin_memory() for _ in range(2)]
        slot = self._slot
        self._slot ^= 1
        cls_arr = (C.c_int * len(self.classes))(*self.classes) if self.classes else None
        ticket = self.ctx.call("aq_detect_submit", C.c_void_p(tiles.data_ptr()), n, h, w, int(bgr), int(src_is_device),
                               self.conf_thres, self.iou_thres, self.max_det, int(self.agnostic), cls_arr,
                               len(self.classes) if self.classes else 0, C.c_void_p(self._dets[slot].data_ptr()),
                               C.c_void_p(self._counts[slot].data_ptr()), stream)
        self._inflight[slot] = tiles  # keeps the source alive while the copy stream may still read it
        return slot, n, ticket

    # ------------------------------------------------------------------ JPEG tiles
    def jpeg_staging(self, n, coef_count, min_entries=0):
        """One of two alternating pinned staging areas for entropy-decoded tiles in packed form: numpy views
        (block_desc uint32 (n, coef_count / 64), entries uint32 (capacity >= min_entries; grown on demand, which drains
        the pipeline first), base uint64 (n), quant uint16 (n, 192)).  Fill them (see `pack_jpeg_tiles`), then call `detect_submit_jpeg`.  An area is reused every
        second call, i.e. after its batch has been waited for in a two-deep pipeline."""
        key = ("jpeg", coef_count)
        if self._pin.get(key) is not None and self._pin[key][0][1].numel() < min_entries:
            self.ctx.sync()  # copies out of the old staging areas may still be in flight
            self._pin[key] = None
        if self._pin.get(key) is None:
            nb = coef_count // 64
            cap = max(min_entries + min_entries // 4, self.batch * coef_count // 8, 1)
            self._pin[key] = [(torch.empty(self.batch, nb, dtype=torch.int32).pin_memory(),
                               torch.empty(cap, dtype=torch.int32).pin_memory(),
                               torch.empty(self.batch, dtype=torch.int64).pin_memory(),
                               torch.empty(self.batch, 192, dtype=torch.uint16).pin_memory()) for _ in range(2)]
            self._pin_i[key] = 0
        i = self._pin_i[key]
        self._pin_i[key] = i ^ 1
        d, e, b, q = self._pin[key][i]
        return d[:n].numpy().view(np.uint32), e.numpy().view(np.uint32), b[:n].numpy().view(np.uint64), q[:n].numpy()

    def decode_jpeg_batch(self, block_desc, entries, base, total_entries, quant, width, height, ncomp, hs, bgr=True, stream=None) -> torch.Tensor:
        """Device half of the hybrid JPEG decoder (aq_jpeg_reconstruct_packed): packed coefficients + tables of n
        equally-shaped tiles (host arrays) -> uint8 (n, height, width, 3) cuda tensor, bit-identical to cv2.imread of the
        same files.  H2D on the library's copy stream, kernels on the context's stream (not synchronised): consume the
        result with `detect_submit(..., src_is_device=True)` or call `ctx.sync()` before reading it elsewhere."""
        n = block_desc.shape[0]
        assert block_desc.dtype == np.uint32 and entries.dtype == np.uint32 and base.dtype == np.uint64 and quant.shape == (n, 192)
        key = (height, width)
        if self._jpeg_out.get(key) is None:
            self._jpeg_out[key] = torch.empty(self.batch, height, width, 3, dtype=torch.uint8, device=f"cuda:{self.ctx.device}")
            torch.cuda.synchronize(self.ctx.device)
        out = self._jpeg_out[key][:n]
        self.ctx.call("aq_jpeg_reconstruct_packed", C.c_void_p(block_desc.ctypes.data), C.c_void_p(entries.ctypes.data),
                      C.c_void_p(base.ctypes.data), int(total_entries), C.c_void_p(quant.ctypes.data), n, width, height, ncomp, hs,
                      int(bgr), C.c_void_p(out.data_ptr()), stream)
        return out

    def detect_submit_jpeg(self, block_desc, entries, base, total_entries, quant, width, height, ncomp, hs, stream=None):
        """JPEG tiles -> detections without the decoded pixels ever visiting the host: reconstruct on the device, then the
        usual batch (letterbox, conv stack, decode, NMS).  Returns a ticket for `detect_wait`."""
        dev = self.decode_jpeg_batch(block_desc, entries, base, total_entries, quant, width, height, ncomp, hs, bgr=True, stream=stream)
        return self.detect_submit(dev, bgr=True, src_is_device=True, stream=stream)

    def jpeg_scan_staging(self, n, min_bytes=0):
        """One of two alternating pinned staging areas for the all-device decoder: numpy views (scans uint8 (capacity >=
        min_bytes), off uint64 (n), length uint32 (n), tables uint8 (n, AQ_JPEG_TABLE_BYTES), quant uint16 (n, 192),
        status int32 (n)).  Fill with `pack_jpeg_scans`, submit with `detect_submit_jpeg_scans`."""
        key = "jpegscan"
        if self._pin.get(key) is not None and self._pin[key][0][0].numel() < min_bytes:
            self.ctx.sync()
            self._pin[key] = None
        if self._pin.get(key) is None:
            cap = max(min_bytes + min_bytes // 2, self.batch * (256 << 10))
            self._pin[key] = [(torch.empty(cap, dtype=torch.uint8).pin_memory(), torch.empty(self.batch, dtype=torch.int64).pin_memory(),
                               torch.empty(self.batch, dtype=torch.int32).pin_memory(),
                               torch.empty(self.batch, _lib.AQ_JPEG_TABLE_BYTES, dtype=torch.uint8).pin_memory(),
                               torch.empty(self.batch, 192, dtype=torch.uint16).pin_memory(),
                               torch.zeros(self.batch, dtype=torch.int32).pin_memory()) for _ in range(2)]
            self._pin_i[key] = 0
        i = self._pin_i[key]
        self._pin_i[key] = i ^ 1
        sc, off, ln, tab, q, stt = self._pin[key][i]
        return (sc.numpy(), off[:n].numpy().view(np.uint64), ln[:n].numpy().view(np.uint32), tab[:n].numpy(), q[:n].numpy(),
                stt[:n].numpy())

    def detect_submit_jpeg_scans(self, scans, off, length, total_bytes, tables, quant, status, width, height, ncomp, hs, stream=None):
        """JPEG tiles -> detections with the whole decoder on the device (aq_jpeg_decode_device: parallel Huffman decoding,
        IDCT, upsampling, colour), then the usual batch.  `status` (pinned int32, n) is filled by the time the ticket has
        been waited for: non-zero = that tile's scan was corrupt."""
        n = off.shape[0]
        key = (height, width)
        if self._jpeg_out.get(key) is None:
            self._jpeg_out[key] = torch.empty(self.batch, height, width, 3, dtype=torch.uint8, device=f"cuda:{self.ctx.device}")
            torch.cuda.synchronize(self.ctx.device)
        out = self._jpeg_out[key][:n]
        self.ctx.call("aq_jpeg_decode_device", C.c_void_p(scans.ctypes.data), C.c_void_p(off.ctypes.data), C.c_void_p(length.ctypes.data),
                      int(total_bytes), C.c_void_p(tables.ctypes.data), C.c_void_p(quant.ctypes.data), n, width, height, ncomp, hs, 1,
                      C.c_void_p(out.data_ptr()), C.c_void_p(status.ctypes.data), stream)
        return self.detect_submit(out, bgr=True, src_is_device=True, stream=stream)

    def pinned_input(self, n, h, w) -> torch.Tensor:
        """One of two alternating pinned uint8 (n, h, w, 3) host buffers for `detect_submit` (a buffer is reused
        every second call, i.e. after its batch has been waited for in a two-deep pipeline)."""
        key = (h, w)
        if self._pin.get(key) is None:
            self._pin[key] = [torch.empty(self.batch, h, w, 3, dtype=torch.uint8).pin_memory() for _ in range(2)]
            self._pin_i[key] = 0
        i = self._pin_i[key]
        self._pin_i[key] = i ^ 1
        return self._pin[key][i][:n]

    def _add_stage_ms(self, ticket: int, n: int):
        """Upstream's three `Speed:` stages of one batch (aq_detect_stage_ms: CUDA events on the compute stream).  The first
        batch of a detector's life is kept apart (`first_stage_ms`): its stages contain one-time work (NMS scratch and output
        buffers are allocated between the events) -- upstream warms the model up before it starts its timers too."""
        ms = (C.c_float * 3)()
        self.ctx.call("aq_detect_stage_ms", ticket, ms)
        if self.first_stage_tiles == 0:
            self.first_stage_ms, self.first_stage_tiles = [ms[0], ms[1], ms[2]], n
            return
        for k in range(3):
            self.stage_ms[k] += ms[k]
        self.stage_tiles += n

    def detect_wait(self, ticket, stream=None) -> Detections:
        """Results of one submitted batch (views of the pinned slot: copy them before the slot's next submit)."""
        slot, n, t = ticket
        self.ctx.call("aq_detect_wait", t)
        self._add_stage_ms(t, n)
        self._inflight[slot] = None
        return Detections(self._dets[slot][:n].numpy(), self._counts[slot][:n].numpy())

    def detect(self, tiles, bgr=True, src_is_device=False) -> Detections:
        return self.detect_wait(self.detect_submit(tiles, bgr, src_is_device))

    def detect_windows(self, raster: torch.Tensor, tile: int, win0: int, n: int, bgr=True) -> Detections:
        """Tiling fused into inference: windows [win0, win0 + n) of the grid of full `tile` x `tile` windows of a
        device-resident (H, W, 3) uint8 raster, enumerated like the reference's tiler (x outer, y inner)
        [REF src/load_data/tile_tifs.py:33-47].  Results are copies (not views of the pinned slot)."""
        assert raster.is_cuda and raster.dtype == torch.uint8 and raster.is_contiguous() and raster.shape[2] == 3
        if self._dets is None:
            self._dets = [torch.empty(self.batch, self.max_det, 6, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._counts = [torch.empty(self.batch, dtype=torch.int32).pin_memory() for _ in range(2)]
        cls_arr = (C.c_int * len(self.classes))(*self.classes) if self.classes else None
        torch.cuda.current_stream(raster.device).synchronize()  # the raster was produced on torch's stream
        slot = self.ctx.call("aq_detect_windows", C.c_void_p(raster.data_ptr()), raster.shape[0], raster.shape[1], tile, tile, win0, n,
                      int(bgr), self.conf_thres, self.iou_thres, self.max_det, int(self.agnostic), cls_arr,
                      len(self.classes) if self.classes else 0, C.c_void_p(self._dets[0].data_ptr()),
                      C.c_void_p(self._counts[0].data_ptr()), None)
        self._add_stage_ms(slot, n)
        return Detections(self._dets[0][:n].numpy().copy(), self._counts[0][:n].numpy().copy())

    def detect_window_list(self, raster: torch.Tensor, tile: int, win_idx: Sequence[int], bgr=True) -> Detections:
        """detect_windows for an arbitrary list of window indices (<= batch of them): how --skip-blank keeps the windows
        src/utils.py::is_blank calls blank away from the network [REF src/utils.py:325-350; the reference does not
        download such images, REF src/load_data/download_french_data.py:58-60]."""
        assert raster.is_cuda and raster.dtype == torch.uint8 and raster.is_contiguous() and raster.shape[2] == 3
        n = len(win_idx)
        if self._dets is None:
            self._dets = [torch.empty(self.batch, self.max_det, 6, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._counts = [torch.empty(self.batch, dtype=torch.int32).pin_memory() for _ in range(2)]
        cls_arr = (C.c_int * len(self.classes))(*self.classes) if self.classes else None
        idx = (C.c_int * n)(*[int(i) for i in win_idx])
        torch.cuda.current_stream(raster.device).synchronize()  # the raster was produced on torch's stream
        slot = self.ctx.call("aq_detect_window_list", C.c_void_p(raster.data_ptr()), raster.shape[0], raster.shape[1], tile, tile, idx, n,
                             int(bgr), self.conf_thres, self.iou_thres, self.max_det, int(self.agnostic), cls_arr,
                             len(self.classes) if self.classes else 0, C.c_void_p(self._dets[0].data_ptr()),
                             C.c_void_p(self._counts[0].data_ptr()), None)
        self._add_stage_ms(slot, n)
        return Detections(self._dets[0][:n].numpy().copy(), self._counts[0][:n].numpy().copy())

    def close(self):
        self.ctx.close()


def window_grid(raster_h: int, raster_w: int, tile: int) -> List[Tuple[int, int]]:
    """(x, y) origins of the full windows in the reference tiler's order (x outer, y inner); the reference names the
    tile `<tif>_<x>_<y>` [REF src/load_data/tile_tifs.py:33-41]."""
    return [(x, y) for x in range(0, raster_w - tile + 1, tile) for y in range(0, raster_h - tile + 1, tile)]


def tile_stats(ctx: Context, src: torch.Tensor, bgr=True, tile: int = 0, win0: int = 0, n: int = 0) -> np.ndarray:
    """Blank-tile measurements on the device (aq_tile_stats): src = (n, h, w, 3) uint8 cuda tiles, or -- with `tile` > 0 --
    an (H, W, 3) raster whose windows [win0, win0 + n) are measured.  Returns (n, 4) int32: L min, L max, blank rows,
    blank columns."""
    assert src.is_cuda and src.dtype == torch.uint8 and src.is_contiguous()
    torch.cuda.current_stream(src.device).synchronize()
    if tile > 0:
        out = np.zeros((n, 4), np.int32)
        ctx.call("aq_tile_stats", C.c_void_p(src.data_ptr()), n, tile, tile, int(bgr), src.shape[0], src.shape[1], win0,
                 out.ctypes.data, None)
    else:
        n = src.shape[0]
        out = np.zeros((n, 4), np.int32)
        ctx.call("aq_tile_stats", C.c_void_p(src.data_ptr()), n, src.shape[1], src.shape[2], int(bgr), 0, 0, 0,
                 out.ctypes.data, None)
    return out


class PackedJpeg:
    """One entropy-decoded tile in packed form (host): geometry, block descriptors, entries, tables."""
    __slots__ = ("info", "block_desc", "entries", "quant", "key")


def entropy_decode_file(blob) -> PackedJpeg:
    """Host half for one file (runs on an io thread; the C parser releases the GIL)."""
    from . import jpeg
    t = PackedJpeg()
    t.info = jpeg.probe(blob)
    t.block_desc = np.empty(t.info.coef_count // 64, np.uint32)
    ent = np.empty(t.info.coef_count, np.uint32)
    t.quant = np.empty(192, np.uint16)
    n = jpeg.entropy_decode_packed(blob, t.block_desc, ent, t.quant)
    t.entries = ent[:n]
    t.key = ("jpeg", t.info.height, t.info.width, t.info.ncomp, t.info.hs)
    return t


def pack_jpeg_tiles(tiles_: Sequence[PackedJpeg], block_desc, entries, base, quant, pool=None) -> int:
    """Concatenate the tiles' entry lists into one staging area (see TileDetector.jpeg_staging); returns the total.
    `pool` (a ThreadPoolExecutor) spreads the copies over the io threads (numpy copies release the GIL)."""
    sizes = [t.entries.size for t in tiles_]
    need = sum(sizes)
    if need > entries.size:
        raise _lib.AquaError(f"pack_jpeg_tiles: {need} entries do not fit the staging area of {entries.size} "
                             f"(pass min_entries to TileDetector.jpeg_staging)")
    starts = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64) if sizes else np.zeros(0, np.int64)

    def one(i):
        t = tiles_[i]
        block_desc[i] = t.block_desc
        quant[i] = t.quant
        base[i] = starts[i]
        entries[starts[i]:starts[i] + sizes[i]] = t.entries

    if pool is None:
        for i in range(len(tiles_)):
            one(i)
    else:
        list(pool.map(one, range(len(tiles_))))
    return need


class ScanJpeg:
    """One tile prepared for the all-device decoder (host): geometry, unstuffed scan, tables."""
    __slots__ = ("info", "scan", "scan_bytes", "tables", "quant", "key")


def prepare_jpeg_file(blob) -> ScanJpeg:
    """Host side of the all-device decoder for one file: marker walk + byte unstuffing only (io thread, GIL released)."""
    from . import jpeg
    t = ScanJpeg()
    t.info, t.scan, t.scan_bytes, t.tables, t.quant = jpeg.scan_prepare(blob)
    t.key = ("jpegscan", t.info.height, t.info.width, t.info.ncomp, t.info.hs)
    return t


class RawJpeg:
    """A JPEG file's bytes + geometry, probed but not yet prepared (the batch call does that: prepare_jpeg_batch)."""
    __slots__ = ("blob", "info", "key")

    def __init__(self, blob, info):
        self.blob, self.info = blob, info
        self.key = ("jpegscan", info.height, info.width, info.ncomp, info.hs)


def decode_jpeg_raster(ctx: Context, blob) -> torch.Tensor:
    """One (large) JPEG file decoded entirely on the device into a uint8 (H, W, 3) BGR cuda tensor -- the raster
    `TileDetector.detect_windows` cuts its windows from, so a .jpeg raster never exists as pixels on the host.
    Synchronises.  Raises jpeg.JpegUnsupported for variants without an all-device path."""
    from . import jpeg
    info, scan, nbytes, tables, quant = jpeg.scan_prepare(blob)
    out = torch.empty(info.height, info.width, 3, dtype=torch.uint8, device=f"cuda:{ctx.device}")
    torch.cuda.synchronize(ctx.device)
    off, length, status = np.zeros(1, np.uint64), np.array([nbytes], np.uint32), np.full(1, -99, np.int32)
    ctx.call("aq_jpeg_decode_device", C.c_void_p(scan.ctypes.data), C.c_void_p(off.ctypes.data), C.c_void_p(length.ctypes.data), scan.size,
             C.c_void_p(tables.ctypes.data), C.c_void_p(quant.ctypes.data), 1, info.width, info.height, info.ncomp, info.hs, 1,
             C.c_void_p(out.data_ptr()), C.c_void_p(status.ctypes.data), None)
    ctx.sync()
    if status[0] != 0:
        raise _lib.AquaError(f"aq_jpeg_decode_device: corrupt JPEG scan (status {int(status[0])})")
    return out


def prepare_jpeg_batch(det: "TileDetector", blobs: Sequence, n_threads: int = 8):
    """Host side of the all-device decoder for a whole batch in ONE library call (aq_jpeg_scan_prepare_batch: marker walk +
    unstuffing on C threads, written straight into a pinned staging area of `det`).  blobs: bytes or uint8 arrays of
    equally-shaped JPEG files.  Returns (staging views, total bytes, info of tile 0) for `detect_submit_jpeg_scans`, or raises
    jpeg.JpegUnsupported / AquaError naming the first offending tile."""
    from . import jpeg
    n = len(blobs)
    arrs = [np.frombuffer(b, np.uint8) if not isinstance(b, np.ndarray) else b for b in blobs]
    lens = np.fromiter((a.size for a in arrs), np.uint64, n)
    need = int(((lens + 16 + 3) // 4 * 4).sum())
    sc, off, ln, tab, q, stt = det.jpeg_scan_staging(n, need)
    ptrs = (C.c_void_p * n)(*[a.ctypes.data for a in arrs])
    infos = (_lib.JpegInfo * n)()
    status = np.zeros(n, np.int32)
    total = C.c_size_t(0)
    rc = _lib.load().aq_jpeg_scan_prepare_batch(ptrs, lens.ctypes.data, n, sc.ctypes.data, sc.size, off.ctypes.data, ln.ctypes.data,
                                                C.byref(total), tab.ctypes.data, q.ctypes.data, infos, status.ctypes.data, n_threads)
    if rc == jpeg.AQ_ERR_UNSUPPORTED:
        raise jpeg.JpegUnsupported(f"tile {int(np.flatnonzero(status)[0])}: JPEG variant without an all-device path")
    if rc < 0:
        raise _lib.AquaError(f"aq_jpeg_scan_prepare_batch failed ({rc}) at tile {int(np.flatnonzero(status)[0]) if status.any() else -1}")
    i0 = infos[0]
    geo = (i0.width, i0.height, i0.ncomp, i0.hs)
    for i in range(1, n):
        if (infos[i].width, infos[i].height, infos[i].ncomp, infos[i].hs) != geo:
            raise _lib.AquaError(f"prepare_jpeg_batch: tile {i} has another geometry than tile 0")
    return (sc, off, ln, tab, q, stt), int(total.value), i0


def pack_jpeg_scans(tiles_: Sequence[ScanJpeg], scans, off, length, tables, quant, pool=None) -> int:
    """Lay the tiles' scans one after another in a staging area (TileDetector.jpeg_scan_staging); returns the extent."""
    sizes = [t.scan.size for t in tiles_]  # multiples of 4, padding included
    need = sum(sizes)
    if need > scans.size:
        raise _lib.AquaError(f"pack_jpeg_scans: {need} bytes do not fit the staging area of {scans.size} (pass min_bytes)")
    starts = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64) if sizes else np.zeros(0, np.int64)

    def one(i):
        t = tiles_[i]
        scans[starts[i]:starts[i] + sizes[i]] = t.scan
        off[i] = starts[i]
        length[i] = t.scan_bytes
        tables[i] = t.tables
        quant[i] = t.quant

    if pool is None:
        for i in range(len(tiles_)):
            one(i)
    else:
        list(pool.map(one, range(len(tiles_))))
    return need


def decode_jpeg_files(ctx: Context, blobs: Sequence[bytes], bgr=True, packed=True, mode=None) -> np.ndarray:
    """Hybrid decode of equally-shaped JPEG files without a planned model (tests, tools): host Huffman decoding +
    aq_jpeg_reconstruct[_packed] -> uint8 (n, h, w, 3) numpy array (BGR = what cv2.imread returns).  The CLI uses the
    pipelined `TileDetector.detect_submit_jpeg` instead."""
    from . import jpeg
    infos = [jpeg.probe(b) for b in blobs]
    i0 = infos[0]
    geo = (i0.width, i0.height, i0.ncomp, i0.hs)
    assert all((i.width, i.height, i.ncomp, i.hs) == geo for i in infos), "files of one batch must share their geometry"
    n = len(blobs)
    out = torch.empty(n, i0.height, i0.width, 3, dtype=torch.uint8, device=f"cuda:{ctx.device}")
    torch.cuda.synchronize(ctx.device)
    if mode == "device":  # entropy decoding on the GPU as well
        tl = [prepare_jpeg_file(b) for b in blobs]
        scans = np.empty(sum(t.scan.size for t in tl), np.uint8)
        off, length = np.empty(n, np.uint64), np.empty(n, np.uint32)
        tables, quant = np.empty((n, _lib.AQ_JPEG_TABLE_BYTES), np.uint8), np.empty((n, 192), np.uint16)
        status = np.full(n, -99, np.int32)
        total = pack_jpeg_scans(tl, scans, off, length, tables, quant)
        ctx.call("aq_jpeg_decode_device", C.c_void_p(scans.ctypes.data), C.c_void_p(off.ctypes.data), C.c_void_p(length.ctypes.data), total,
                 C.c_void_p(tables.ctypes.data), C.c_void_p(quant.ctypes.data), n, i0.width, i0.height, i0.ncomp, i0.hs, int(bgr),
                 C.c_void_p(out.data_ptr()), C.c_void_p(status.ctypes.data), None)
        ctx.sync()
        if status.any():
            raise _lib.AquaError(f"aq_jpeg_decode_device: tile status {status.tolist()}")
    elif packed:
        tl = [entropy_decode_file(b) for b in blobs]
        desc = np.empty((n, i0.coef_count // 64), np.uint32)
        ent = np.empty(max(1, sum(t.entries.size for t in tl)), np.uint32)
        base = np.empty(n, np.uint64)
        quant = np.empty((n, 192), np.uint16)
        total = pack_jpeg_tiles(tl, desc, ent, base, quant)
        ctx.call("aq_jpeg_reconstruct_packed", C.c_void_p(desc.ctypes.data), C.c_void_p(ent.ctypes.data), C.c_void_p(base.ctypes.data),
                 total, C.c_void_p(quant.ctypes.data), n, i0.width, i0.height, i0.ncomp, i0.hs, int(bgr), C.c_void_p(out.data_ptr()), None)
    else:
        coef = np.zeros((n, i0.coef_count), np.int16)
        quant = np.zeros((n, 192), np.uint16)
        for k, b in enumerate(blobs):
            jpeg.entropy_decode(b, coef[k], quant[k])
        ctx.call("aq_jpeg_reconstruct", C.c_void_p(coef.ctypes.data), C.c_void_p(quant.ctypes.data), n, i0.width, i0.height, i0.ncomp,
                 i0.hs, int(bgr), 0, C.c_void_p(out.data_ptr()), None)
    ctx.sync()
    return out.cpu().numpy()


def is_blank(stats_row) -> bool:
    """src/utils.py::is_blank on the measured luminance extrema [REF src/utils.py:339-350]."""
    lo, hi = int(stats_row[0]), int(stats_row[1])
    return (lo, hi) in ((0, 0), (1, 1), (255, 255)) or (lo >= 250 and hi >= 250)


def is_partly_blank(stats_row) -> bool:
    """src/utils.py::is_partly_blank: any fully blank row or column [REF src/utils.py:359-369]."""
    return int(stats_row[2]) + int(stats_row[3]) > 0


def format_label_text(det: np.ndarray, net_hw, hw0, save_conf=True) -> str:
    """Label text of one image via the C formatter (det: (n,6) fp32, descending conf)."""
    lib = _lib.load()
    det = np.ascontiguousarray(det, np.float32)
    cap = 128 * max(len(det), 1) + 16
    out = C.create_string_buffer(cap)
    rc = lib.aq_format_label_text(det.ctypes.data, len(det), int(net_hw[0]), int(net_hw[1]), int(hw0[0]), int(hw0[1]),
                                  int(save_conf), out, cap)
    if rc < 0:
        raise AquaError(f"aq_format_label_text failed ({rc})")
    return out.value.decode()


def write_labels(dets: Detections, net_hw, hw0_list, paths: List[str], save_conf=True, threads=8) -> int:
    """Write one label file per image with >= 1 detection; returns the number of files written."""
    lib = _lib.load()
    n = len(paths)
    d = np.ascontiguousarray(dets.dets, np.float32)
    cnt = np.ascontiguousarray(dets.counts, np.int32)
    h0 = np.asarray([h for h, _ in hw0_list], np.int32)
    w0 = np.asarray([w for _, w in hw0_list], np.int32)
    arr = (C.c_char_p * n)(*[p.encode() for p in paths])
    rc = lib.aq_format_labels(d.ctypes.data, cnt.ctypes.data, n, d.shape[1], int(net_hw[0]), int(net_hw[1]),
                              h0.ctypes.data, w0.ctypes.data, arr, int(save_conf), threads)
    if rc < 0:
        raise AquaError(f"aq_format_labels failed ({rc})")
    return rc


def geocode_label_text(text: str, x_offset: int, y_offset: int, bounds_3857, large_tif_size=6144, im_width=1024, im_height=1024):
    """Rows of one label file -> pixel boxes and EPSG:3857 boxes (host C, aq_geocode_label_text): the per-row arithmetic
    of geocode_all_detections + convert_pix_to_m_bboxes [REF src/process_yolo/geocode_results.py:71-101,158-170].
    Returns a list of dicts in file order."""
    lib = _lib.load()
    cap = text.count("\n") + 1
    rows = (_lib.GeoRow * cap)()
    b = (C.c_double * 4)(*[float(v) for v in bounds_3857])
    n = lib.aq_geocode_label_text(text.encode(), int(x_offset), int(y_offset), b, int(large_tif_size), int(im_width),
                                  int(im_height), C.cast(rows, C.c_void_p), cap)
    if n < 0:
        raise AquaError(f"aq_geocode_label_text failed ({n})")
    return [{f: getattr(rows[i], f) for f, _ in _lib.GeoRow._fields_} for i in range(n)]


def geocode_detections(dets: Detections, net_hw, hw0_list, stems: Sequence[str], download_bounds_3857, large_tif_size=6144,
                       im_width=1024, im_height=1024) -> List[dict]:
    """geocode_all_detections [REF src/process_yolo/geocode_results.py:139-193] straight from a batch's detections (host C,
    aq_geocode_dets): no label files are read, no images opened.  stems[i] = tile stem `<layer><year>_<box>_<xoff>_<yoff>`;
    download_bounds_3857 = (n_boxes, 4) bounds of the download boxes.  Returns one dict per detection in label-file order:
    image (stem + '.jpeg'), pixel box, EPSG:3035 box as xmin_m.., type, year, det_conf, geometry_3857 (xmin, ymin, xmax, ymax)."""
    lib = _lib.load()
    n = len(stems)
    d = np.ascontiguousarray(dets.dets, np.float32)
    cnt = np.ascontiguousarray(dets.counts, np.int32)
    h0 = np.asarray([h for h, _ in hw0_list], np.int32)
    w0 = np.asarray([w for _, w in hw0_list], np.int32)
    arr = (C.c_char_p * n)(*[s.encode() for s in stems])
    b = np.ascontiguousarray(np.asarray(download_bounds_3857, np.float64).reshape(-1, 4))
    cap = int(cnt[:n].clip(min=0).sum())
    rows = (_lib.GeoDet * max(cap, 1))()
    got = lib.aq_geocode_dets(d.ctypes.data, cnt.ctypes.data, n, d.shape[1], int(net_hw[0]), int(net_hw[1]), h0.ctypes.data,
                              w0.ctypes.data, arr, b.ctypes.data, b.shape[0], int(large_tif_size), int(im_width), int(im_height),
                              C.cast(rows, C.c_void_p), cap)
    if got < 0:
        raise AquaError(f"aq_geocode_dets failed ({got}): a stem is not <layer><year>_<box>_<xoff>_<yoff> or names an unknown download box")
    out = []
    for r in rows[:got]:
        name = lib.aq_geo_class_name(r.cls)
        if name is None:
            raise KeyError(r.cls)  # REVERSE_CLASS_MAPPING[bbox[0]] upstream
        out.append(dict(image=stems[r.image] + ".jpeg", xmin=r.xmin, xmax=r.xmax, ymin=r.ymin, ymax=r.ymax, xmin_m=r.xmin_a,
                        xmax_m=r.xmax_a, ymin_m=r.ymin_a, ymax_m=r.ymax_a, type=name.decode(), year=r.year, det_conf=r.conf,
                        geometry_3857=(r.xmin_m, r.ymin_m, r.xmax_m, r.ymax_m)))
    return out


def transform_3857_to_3035(xy) -> np.ndarray:
    """(n, 2) EPSG:3857 -> EPSG:3035 (aq_transform_3857_to_3035; see the header for what pins it)."""
    a = np.ascontiguousarray(np.asarray(xy, np.float64).reshape(-1, 2))
    out = np.empty_like(a)
    rc = _lib.load().aq_transform_3857_to_3035(a.ctypes.data, a.shape[0], out.ctypes.data)
    if rc < 0:
        raise AquaError(f"aq_transform_3857_to_3035 failed ({rc})")
    return out


def cage_areas(square: bool, w: float, h: float, x_border=False, y_border=False):
    """(estimate, variance, min, max) of get_circle_area_from_bbox / get_square_area_from_bbox
    [REF src/process_yolo/calc_net_areas.py:24-82] (host C, aq_cage_areas)."""
    out = (C.c_double * 4)()
    rc = _lib.load().aq_cage_areas(int(bool(square)), float(w), float(h), int(bool(x_border)), int(bool(y_border)), out)
    if rc < 0:
        raise AquaError(f"aq_cage_areas failed ({rc})")
    return tuple(out)


def dbscan_labels(center_points, distance_threshold: float, min_samples: int) -> np.ndarray:
    """get_clusters [REF src/cluster_facilities.py:11-15]: labels of sklearn's DBSCAN(eps, min_samples) on 2-D points
    (host C, aq_dbscan); -1 = noise."""
    pts = np.ascontiguousarray(np.asarray(center_points, np.float64).reshape(-1, 2))
    labels = np.empty(pts.shape[0], np.int32)
    rc = _lib.load().aq_dbscan(pts.ctypes.data, pts.shape[0], float(distance_threshold), int(min_samples), labels.ctypes.data)
    if rc < 0:
        raise AquaError(f"aq_dbscan failed ({rc})")
    return labels
