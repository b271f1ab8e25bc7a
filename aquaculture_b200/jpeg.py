"""Host side of the hybrid JPEG tile decoder (include/aqua_b200.h "JPEG tiles").

Replaces the cv2.imread call of upstream utils/dataloaders.py::LoadImages.__next__ [UPSTREAM-RECALLED] for the tiles
gdal.Translate -of JPEG writes [REF src/load_data/tile_tifs.py:74]: the C parser in the library undoes the Huffman coding
on host threads (ctypes releases the GIL), the GPU does dequantisation, IDCT, chroma upsampling and colour conversion
(`TileDetector.decode_jpeg_batch` in engine.py).  Files the parser does not cover (progressive, CMYK, 4:2:2 ...) raise
`JpegUnsupported`; the caller decides what to do with them -- nothing is decoded silently another way here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

AQ_ERR_UNSUPPORTED = -5


class JpegUnsupported(_lib.AquaError):
    pass


def probe(data: bytes | np.ndarray) -> _lib.JpegInfo:
    """Geometry of a JPEG file (aq_jpeg_probe).  `data`: the file's bytes."""
    lib = _lib.load()
    buf = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else data
    info = _lib.JpegInfo()
    rc = lib.aq_jpeg_probe(buf.ctypes.data, buf.size, C.byref(info))
    if rc == AQ_ERR_UNSUPPORTED:
        raise JpegUnsupported("JPEG variant without a device path (progressive / arithmetic / CMYK / 4:2:2 ...)")
    if rc < 0:
        raise _lib.AquaError(f"aq_jpeg_probe failed ({rc}): not a decodable JPEG file")
    return info


def entropy_decode(data: bytes | np.ndarray, coef: np.ndarray, quant: np.ndarray) -> None:
    """Huffman-decode one file into `coef` (int16, >= info.coef_count values) and `quant` (uint16, 192)."""
    lib = _lib.load()
    buf = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else data
    assert coef.dtype == np.int16 and coef.flags.c_contiguous and quant.dtype == np.uint16 and quant.size >= 192
    rc = lib.aq_jpeg_entropy_decode(buf.ctypes.data, buf.size, coef.ctypes.data, coef.size, quant.ctypes.data)
    if rc == AQ_ERR_UNSUPPORTED:
        raise JpegUnsupported("JPEG variant without a device path")
    if rc < 0:
        raise _lib.AquaError(f"aq_jpeg_entropy_decode failed ({rc}): corrupt or truncated entropy-coded data")


def entropy_decode_packed(data: bytes | np.ndarray, block_desc: np.ndarray, entries: np.ndarray, quant: np.ndarray) -> int:
    """Huffman-decode one file into the packed form (aq_jpeg_entropy_decode_packed): block_desc uint32 (>= coef_count / 64),
    entries uint32 (>= coef_count), quant uint16 (192).  Returns the number of entries written."""
    lib = _lib.load()
    buf = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else data
    assert block_desc.dtype == np.uint32 and entries.dtype == np.uint32 and quant.dtype == np.uint16 and quant.size >= 192
    assert block_desc.flags.c_contiguous and entries.flags.c_contiguous
    n = C.c_size_t(0)
    rc = lib.aq_jpeg_entropy_decode_packed(buf.ctypes.data, buf.size, block_desc.ctypes.data, block_desc.size, entries.ctypes.data,
                                           entries.size, C.byref(n), quant.ctypes.data)
    if rc == AQ_ERR_UNSUPPORTED:
        raise JpegUnsupported("JPEG variant without a device path")
    if rc < 0:
        raise _lib.AquaError(f"aq_jpeg_entropy_decode_packed failed ({rc}): corrupt or truncated entropy-coded data")
    return int(n.value)


def unpack(block_desc: np.ndarray, entries: np.ndarray) -> np.ndarray:
    """Packed form -> dense int16 (blocks * 64) coefficients (tests; the device does this inside the IDCT kernel)."""
    nb = block_desc.size
    coef = np.zeros(nb * 64, np.int16)
    first, cnt = (block_desc >> 7).astype(np.int64), (block_desc & 127).astype(np.int64)
    blk = np.repeat(np.arange(nb, dtype=np.int64), cnt)
    idx = np.repeat(first - np.concatenate([[0], np.cumsum(cnt)[:-1]]), cnt) + np.arange(cnt.sum(), dtype=np.int64)
    e = entries[idx]
    coef[blk * 64 + ((e >> 16) & 63)] = (e & 0xFFFF).astype(np.uint16).view(np.int16)
    return coef


def scan_prepare(data: bytes | np.ndarray):
    """Host side of the all-device decoder (aq_jpeg_scan_prepare): marker walk only.  Returns (info, scan, scan_bytes,
    tables, quant): `scan` = the entropy-coded segment without byte stuffing, zero padded to whole words + 8 bytes (uint8,
    4-byte aligned), `scan_bytes` its unpadded length, `tables` the six Huffman tables in device form, `quant` 192 uint16."""
    lib = _lib.load()
    buf = np.frombuffer(data, np.uint8) if not isinstance(data, np.ndarray) else data
    info = _lib.JpegInfo()
    scan = np.empty((buf.size + 16 + 3) // 4, np.uint32).view(np.uint8)
    tables = np.empty(_lib.AQ_JPEG_TABLE_BYTES // 4, np.uint32).view(np.uint8)
    quant = np.empty(192, np.uint16)
    n = C.c_size_t(0)
    rc = lib.aq_jpeg_scan_prepare(buf.ctypes.data, buf.size, C.byref(info), scan.ctypes.data, scan.size, C.byref(n), tables.ctypes.data,
                                  quant.ctypes.data)
    if rc == AQ_ERR_UNSUPPORTED:
        raise JpegUnsupported("JPEG variant without an all-device path (progressive / arithmetic / CMYK / 4:2:2 / restart intervals)")
    if rc < 0:
        raise _lib.AquaError(f"aq_jpeg_scan_prepare failed ({rc}): not a decodable JPEG file")
    nb = int(n.value)
    return info, scan[:(nb + 3) // 4 * 4 + 8], nb, tables, quant


def debug_parallel_decode(info, scan: np.ndarray, scan_bytes: int, tables: np.ndarray, sub_bits: int = 1024):
    """CPU rehearsal of the device's parallel Huffman schedule (tests): returns (dense int16 coefficients, rounds)."""
    lib = _lib.load()
    coef = np.empty(info.coef_count, np.int16)
    rounds = C.c_int(0)
    rc = lib.aq_jpeg_debug_parallel_decode(scan.ctypes.data, scan_bytes, tables.ctypes.data, C.byref(info), sub_bits, coef.ctypes.data,
                                           coef.size, C.byref(rounds))
    if rc < 0:
        raise _lib.AquaError(f"aq_jpeg_debug_parallel_decode failed ({rc})")
    return coef, rounds.value
