"""aq_tile_stats_static: the allocation-free variant of aq_tile_stats (column sums in a file-scope static __device__ array
of the library, results into the caller's device buffer).  It has no counterpart in the reference;
the checker is the existing entry point aq_tile_stats, whose own parity with the reference's
is_blank / is_partly_blank is pinned elsewhere (tests/test_gpu_kernels.py, tests/golden/).  Integer work: bit-exact.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from aquaculture_b200 import engine

pytestmark = pytest.mark.gpu

MAX_COLS = 131072  # AQ_TILE_STATS_STATIC_MAX_COLS (include/aqua_b200.h)


def _tiles(n, h, w, seed):
    rng = np.random.default_rng(seed)
    t = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    t[0] = 255                        # a blank tile
    t[1 % n, : h // 3] = 255          # blank rows
    t[2 % n, :, w // 2:] = 252        # blank columns (mean >= 250)
    t[3 % n] = rng.integers(240, 256, (h, w, 3), dtype=np.uint8)  # near the threshold
    return torch.from_numpy(t).cuda()


def _reference(ctx, src, raster=None, tile=0, win0=0, n=0):
    if raster is not None:
        return engine.tile_stats(ctx, raster, bgr=True, tile=tile, win0=win0, n=n)
    return engine.tile_stats(ctx, src, bgr=True)


def _static(ctx, src, n, h, w, out, raster_hw=(0, 0), win0=0):
    torch.cuda.synchronize()  # inputs were produced on torch's stream; the library works on the context's own
    ctx.call("aq_tile_stats_static", C.c_void_p(src.data_ptr()), n, h, w, 1, raster_hw[0], raster_hw[1], win0,
             C.c_void_p(out.data_ptr()), None)
    ctx.sync()
    return out[:n].cpu().numpy()


def test_tile_stats_static_matches_tile_stats_on_two_calls_in_a_row():
    ctx = engine.Context(0)
    try:
        out = torch.full((64, 4), -7, dtype=torch.int32, device="cuda")  # the caller's output buffer, allocated up front
        # call 1: 24 tiles of 640 x 640 (vectorised kernel); call 2 right behind it: fewer, narrower tiles with other
        # contents (scalar kernel) -- the shared array still holds call 1's column sums where call 2 does not clear them
        a = _tiles(24, 640, 640, seed=1)
        b = _tiles(7, 333, 250, seed=2)
        torch.cuda.synchronize()
        got_a = _static(ctx, a, 24, 640, 640, out).copy()
        got_b = _static(ctx, b, 7, 333, 250, out).copy()
        ref_a, ref_b = _reference(ctx, a), _reference(ctx, b)
        assert np.array_equal(got_a, np.asarray(ref_a).reshape(24, 4))
        assert np.array_equal(got_b, np.asarray(ref_b).reshape(7, 4))
        assert engine.is_blank(got_a[0]) and not engine.is_blank(got_a[5])
        # the same first call again (the array now holds call 2's sums), and windows of a raster
        assert np.array_equal(_static(ctx, a, 24, 640, 640, out), got_a)
        raster = _tiles(1, 1024, 1536, seed=3)[0].contiguous()
        got_w = _static(ctx, raster, 6, 512, 512, out, raster_hw=(1024, 1536), win0=0).copy()
        ref_w = _reference(ctx, None, raster=raster, tile=512, win0=0, n=6)
        assert np.array_equal(got_w, np.asarray(ref_w).reshape(6, 4))
    finally:
        ctx.close()


def test_tile_stats_static_rejects_sizes_beyond_the_static_scratch():
    ctx = engine.Context(0)
    try:
        out = torch.zeros((256, 4), dtype=torch.int32, device="cuda")
        n, w = 129, 1024  # 132 096 columns > 131 072
        src = torch.zeros((n, 8, w, 3), dtype=torch.uint8, device="cuda")
        with pytest.raises(engine.AquaError, match="exceed the static scratch"):
            ctx.call("aq_tile_stats_static", C.c_void_p(src.data_ptr()), n, 8, w, 1, 0, 0, 0, C.c_void_p(out.data_ptr()), None)
        # exactly at the limit is accepted and agrees with the allocating entry point
        n = MAX_COLS // w
        got = _static(ctx, src[:n], n, 8, w, out)
        assert np.array_equal(got, np.asarray(engine.tile_stats(ctx, src[:n], bgr=True)).reshape(n, 4))
    finally:
        ctx.close()
