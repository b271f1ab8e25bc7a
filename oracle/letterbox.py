"""ORACLE (test infrastructure): letterbox + to-tensor of the reference's input path.

Restates upstream `utils/augmentations.py::letterbox` and the `LoadImages.__next__` /
`detect.py::run` pre-process block [UPSTREAM-RECALLED, SURVEY.md §8 a-2..a-4]; the reference
invokes them through `yolov5/detect.py` [REF README.md:77] on 1024x1024 tiles
[REF src/utils.py:18-19].  Pinned by cv2 itself (`cv2.resize`, `cv2.copyMakeBorder` are the
genuine third-party kernels): `resize_bilinear_u8` below is an integer restatement of cv2's
INTER_LINEAR uint8 path and tests/test_oracle.py checks it bit-for-bit against cv2.
"""
from __future__ import annotations

import numpy as np


def letterbox_geometry(h0: int, w0: int, new_shape=(640, 640), auto=True, scale_fill=False,
                       scaleup=True, stride=32):
    """Returns (new_unpad_w, new_unpad_h, top, bottom, left, right) exactly as upstream computes."""
    if isinstance(new_shape, int):
        new_shape = (new_shape, new_shape)
    r = min(new_shape[0] / h0, new_shape[1] / w0)
    if not scaleup:
        r = min(r, 1.0)
    new_unpad = int(round(w0 * r)), int(round(h0 * r))
    dw, dh = new_shape[1] - new_unpad[0], new_shape[0] - new_unpad[1]
    if auto:
        dw, dh = np.mod(dw, stride), np.mod(dh, stride)
    elif scale_fill:
        dw, dh = 0.0, 0.0
        new_unpad = (new_shape[1], new_shape[0])
    dw /= 2
    dh /= 2
    top, bottom = int(round(dh - 0.1)), int(round(dh + 0.1))
    left, right = int(round(dw - 0.1)), int(round(dw + 0.1))
    return new_unpad[0], new_unpad[1], top, bottom, left, right


def _coeffs(dst: int, src: int, vertical: bool):
    """cv2 INTER_LINEAR uint8: source indices and 11-bit fixed-point weights per destination sample.
    Horizontally cv2 zeroes the fraction where the 2-tap window leaves the image; vertically it keeps
    the weights and clips the ROW INDICES instead (both taps then read the border row)."""
    scale = 1.0 / (np.float64(dst) / np.float64(src))   # cv2: scale_x = 1. / inv_scale_x
    d = np.arange(dst)
    fx = ((d + 0.5) * scale - 0.5).astype(np.float32)
    i0 = np.floor(fx).astype(np.int32)
    f = (fx - i0).astype(np.float32)
    if not vertical:
        lo = i0 < 0
        f[lo] = 0.0
        i0[lo] = 0
        hi = i0 >= src - 1
        f[hi] = 0.0
        i0[hi] = src - 1
    i1 = np.clip(i0 + 1, 0, src - 1)
    i0 = np.clip(i0, 0, src - 1)
    a1 = np.rint(f * np.float32(2048)).astype(np.int32)
    a0 = np.rint((np.float32(1.0) - f) * np.float32(2048)).astype(np.int32)
    return i0, i1, a0, a1


def resize_bilinear_u8(im: np.ndarray, new_w: int, new_h: int) -> np.ndarray:
    """Integer emulation of cv2.resize(im, (new_w, new_h), interpolation=cv2.INTER_LINEAR) for uint8 HWC."""
    h, w = im.shape[:2]
    x0, x1, ax0, ax1 = _coeffs(new_w, w, vertical=False)
    y0, y1, ay0, ay1 = _coeffs(new_h, h, vertical=True)
    src = im.astype(np.int32)
    rows = src[:, x0, :] * ax0[None, :, None] + src[:, x1, :] * ax1[None, :, None]   # (h, new_w, c)
    s0 = rows[y0] >> 4
    s1 = rows[y1] >> 4
    out = (((ay0[:, None, None] * s0) >> 16) + ((ay1[:, None, None] * s1) >> 16) + 2) >> 2
    return out.astype(np.uint8)


def letterbox(im: np.ndarray, new_shape=(640, 640), color=(114, 114, 114), auto=True, scaleup=True,
              stride=32, use_cv2=True) -> np.ndarray:
    """uint8 HWC in -> resized + padded uint8 HWC out."""
    h0, w0 = im.shape[:2]
    nw, nh, top, bottom, left, right = letterbox_geometry(h0, w0, new_shape, auto, False, scaleup, stride)
    if (w0, h0) != (nw, nh):
        if use_cv2:
            import cv2
            im = cv2.resize(im, (nw, nh), interpolation=cv2.INTER_LINEAR)
        else:
            im = resize_bilinear_u8(im, nw, nh)
    out = np.empty((nh + top + bottom, nw + left + right, im.shape[2]), np.uint8)
    out[...] = np.asarray(color, np.uint8)
    out[top:top + nh, left:left + nw] = im
    return out


def to_tensor(im_bgr: np.ndarray) -> np.ndarray:
    """LoadImages: HWC BGR -> CHW RGB contiguous; detect.py: float()/255 (fp32)."""
    chw = np.ascontiguousarray(im_bgr.transpose(2, 0, 1)[::-1])
    return chw.astype(np.float32) / np.float32(255.0)
