"""TEST INFRASTRUCTURE ONLY (imported by tests/, never by the product path).

CPU restatement (numpy, integer arithmetic) of the reconstruction half of JPEG decoding as cv2.imread performs it
inside upstream utils/dataloaders.py::LoadImages.__next__ [UPSTREAM-RECALLED] on the tiles written by
gdal.Translate -of JPEG [REF src/load_data/tile_tifs.py:74].  The arithmetic lives in a third-party dependency
absent from /root/reference: libjpeg-turbo (bundled in opencv-python 4.7.0.72 [REF requirements.txt:151]; this
container's cv2 4.13 bundles libjpeg-turbo 3.1.2).  Restated from the published algorithms:
  * inverse DCT: Loeffler-Ligtenberg-Moschytz 8-point IDCT in 13-bit fixed point with 2 guard bits between the
    column and the row pass (libjpeg's default "islow" method), output clamp +128 -> [0, 255];
  * chroma upsampling 4:2:0: 3/4-1/4 triangle filter per axis ("fancy upsampling", the default), rounding terms
    8 / 7 alternating, edge samples replicated;
  * colour: JFIF YCbCr -> RGB with 16-bit fixed-point constants.
Pinned by the genuine library: tests compare decode() with cv2.imdecode bit for bit on synthetic files and on the
committed vectors of tests/golden/jpeg_ref.npz (tools/make_golden_jpeg.py).

The entropy-decoded input (coefficients, tables, geometry) comes from the product's host parser
(aq_jpeg_entropy_decode); identity of the final image with cv2 pins that parser as well.
"""
from __future__ import annotations

import numpy as np

CONST_BITS, PASS1_BITS = 13, 2
F = dict(f0298=2446, f0390=3196, f0541=4433, f0765=6270, f0899=7373, f1175=9633, f1501=12299, f1847=15137, f1961=16069,
         f2053=16819, f2562=20995, f3072=25172)


def _idct8(s):
    """One 1-D pass over axis -1 of `s` (int64 [..., 8]); returns the eight outputs before descaling."""
    s0, s1, s2, s3, s4, s5, s6, s7 = (s[..., i] for i in range(8))
    z1 = (s2 + s6) * F["f0541"]
    tmp2 = z1 - s6 * F["f1847"]
    tmp3 = z1 + s2 * F["f0765"]
    tmp0 = (s0 + s4) << CONST_BITS
    tmp1 = (s0 - s4) << CONST_BITS
    tmp10, tmp13, tmp11, tmp12 = tmp0 + tmp3, tmp0 - tmp3, tmp1 + tmp2, tmp1 - tmp2
    t0, t1, t2, t3 = s7, s5, s3, s1
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * F["f1175"]
    t0, t1, t2, t3 = t0 * F["f0298"], t1 * F["f2053"], t2 * F["f3072"], t3 * F["f1501"]
    z1, z2 = -z1 * F["f0899"], -z2 * F["f2562"]
    z3, z4 = -z3 * F["f1961"] + z5, -z4 * F["f0390"] + z5
    t0, t1, t2, t3 = t0 + z1 + z3, t1 + z2 + z4, t2 + z2 + z3, t3 + z1 + z4
    return np.stack([tmp10 + t3, tmp11 + t2, tmp12 + t1, tmp13 + t0, tmp13 - t0, tmp12 - t1, tmp11 - t2, tmp10 - t3], -1)


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def idct_blocks(coef: np.ndarray, quant: np.ndarray) -> np.ndarray:
    """coef int16 [nblocks, 64] natural order, quant uint16 [64] -> uint8 [nblocks, 8, 8]."""
    x = coef.astype(np.int64).reshape(-1, 8, 8) * quant.astype(np.int64).reshape(1, 8, 8)
    ws = _descale(_idct8(x.transpose(0, 2, 1)), CONST_BITS - PASS1_BITS).transpose(0, 2, 1)  # columns first
    o = _descale(_idct8(ws), CONST_BITS + PASS1_BITS + 3)
    return np.clip(o + 128, 0, 255).astype(np.uint8)


def plane_from_blocks(blocks: np.ndarray, bh: int, bw: int) -> np.ndarray:
    return blocks.reshape(bh, bw, 8, 8).transpose(0, 2, 1, 3).reshape(bh * 8, bw * 8)


def upsample_h2v2(c: np.ndarray, height: int, width: int) -> np.ndarray:
    """Triangle-filter 2x upsampling of a chroma plane whose real extent is ceil(height/2) x ceil(width/2)."""
    ch, cw = (height + 1) // 2, (width + 1) // 2
    c = c[:ch, :cw].astype(np.int32)
    if cw <= 2:  # libjpeg only applies the triangle filter to components more than two samples wide: box replication
        return np.repeat(np.repeat(c, 2, 0), 2, 1)[:height, :width]
    up = np.concatenate([c[:1], c[:-1]], 0)    # row above (edge replicated)
    dn = np.concatenate([c[1:], c[-1:]], 0)    # row below
    out = np.empty((2 * ch, 2 * cw), np.int32)
    for v, far in ((0, up), (1, dn)):
        s = 3 * c + far
        left = np.concatenate([s[:, :1], s[:, :-1]], 1)
        right = np.concatenate([s[:, 1:], s[:, -1:]], 1)
        out[v::2, 0::2] = (3 * s + left + 8) >> 4
        out[v::2, 1::2] = (3 * s + right + 7) >> 4
    return out[:height, :width]


def ycc_to_bgr(y: np.ndarray, cb: np.ndarray, cr: np.ndarray) -> np.ndarray:
    y = y.astype(np.int32)
    u, v = cb.astype(np.int32) - 128, cr.astype(np.int32) - 128
    r = y + ((91881 * v + 32768) >> 16)
    g = y + ((-22554 * u + (-46802 * v + 32768)) >> 16)
    b = y + ((116130 * u + 32768) >> 16)
    return np.clip(np.stack([b, g, r], -1), 0, 255).astype(np.uint8)


def reconstruct(coef: np.ndarray, quant: np.ndarray, width: int, height: int, ncomp: int, hs: int) -> np.ndarray:
    """coef / quant in the layout aq_jpeg_entropy_decode writes -> (height, width, 3) uint8 BGR (= cv2.imread)."""
    mx, my = -(-width // (8 * hs)), -(-height // (8 * hs))
    lb = mx * hs * my * hs
    quant = quant.reshape(3, 64)
    coef = coef.reshape(-1, 64)
    yp = plane_from_blocks(idct_blocks(coef[:lb], quant[0]), my * hs, mx * hs)[:height, :width]
    if ncomp == 1:
        return np.repeat(yp[:, :, None], 3, 2)
    cbp = plane_from_blocks(idct_blocks(coef[lb:lb + mx * my], quant[1]), my, mx)
    crp = plane_from_blocks(idct_blocks(coef[lb + mx * my:lb + 2 * mx * my], quant[2]), my, mx)
    if hs == 2:
        cbp, crp = upsample_h2v2(cbp, height, width), upsample_h2v2(crp, height, width)
    else:
        cbp, crp = cbp[:height, :width], crp[:height, :width]
    return ycc_to_bgr(yp, cbp, crp)
