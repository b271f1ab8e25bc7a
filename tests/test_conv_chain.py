"""EPI_CHAIN (aq_plan_conv_chain): a 1x1 conv evaluated behind its producer inside the same kernel, on the fp16 groups the
producer's epilogue stages in shared memory.  Both outputs must equal -- BIT FOR BIT -- what two separate launches write
(the chained conv reads exactly the fp16-rounded activations a second launch would read back from memory; only the
accumulation order inside an MMA could differ, and it does not: same K order), and match fp32 torch within fp16 tolerance."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from aquaculture_b200 import engine

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = engine.Context(0)
    yield c
    c.close()


def run_chain_case(ctx, B, H, W, c_in, c1, c_in2, c2, k=1, store_first=True, up_c=0, dst_ct=None, dst2_ct=None, dst2_off=0, mblk=0, seed=0,
                   stride=1, mode=0):
    g = torch.Generator().manual_seed(seed)
    dst_ct, dst2_ct = dst_ct or c1, dst2_ct or c2
    x = torch.randn(B, H, W, c_in - up_c, generator=g).half()
    xu = torch.randn(B, H // 2, W // 2, up_c, generator=g).half() if up_c else None
    w1 = (torch.randn(c1, c_in, k, k, generator=g) / math.sqrt(c_in * k * k)).half().float()
    b1 = torch.randn(c1, generator=g) * 0.5
    w2 = (torch.randn(c2, c_in2, 1, 1, generator=g) / math.sqrt(c_in2)).half().float()
    b2 = torch.randn(c2, generator=g) * 0.5
    outs = {}
    for chained in (False, True):
        ctx.plan_begin(B, 64, 64, 3)
        bs = ctx.buffer(H, W, c_in - up_c)
        bd = ctx.buffer(H // stride, W // stride, dst_ct)
        bd2 = ctx.buffer(H // stride, W // stride, dst2_ct)
        bu = ctx.buffer(H // 2, W // 2, up_c) if up_c else None
        up = (bu, 0, up_c) if up_c else None
        if chained:
            ctx.conv_chain((bs, 0), (bd, 0), w1, b1, k, stride, 2, (bd2, dst2_off), w2, b2, 2, store_first=store_first, up=up, mblk=mblk, mode=mode)
        else:
            ctx.conv((bs, 0), (bd, 0), w1, b1, k, stride, 2, up=up, mblk=mblk, mode=mode)
            ctx.conv((bd, 0), (bd2, dst2_off), w2, b2, 1, 1, 2)
        ctx.plan_end()
        ctx.upload(bs, x.numpy())
        if up_c:
            ctx.upload(bu, xu.numpy())
        ctx.upload(bd, np.full((B, H // stride, W // stride, dst_ct), 7.0, np.float16))
        ctx.upload(bd2, np.full((B, H // stride, W // stride, dst2_ct), 5.0, np.float16))
        ctx.forward()
        ctx.sync()
        outs[chained] = (ctx.download(bd), ctx.download(bd2), ctx.call("aq_plan_num_launches"))
    (y1, y2, n_sep), (z1, z2, n_ch) = outs[False], outs[True]
    assert n_sep == 2 and n_ch == 1
    def where(a, b):
        idx = np.argwhere(a != b)
        return (f"{len(idx)} values; images {sorted(set(idx[:, 0].tolist()))[:6]}, rows {idx[:, 1].min()}..{idx[:, 1].max()}, cols "
                f"{idx[:, 2].min()}..{idx[:, 2].max()}, channels {sorted(set((idx[:, 3] // 16).tolist()))} (x16); first {idx[:3].tolist()}: "
                f"{[float(a[tuple(i)]) for i in idx[:3]]} vs {[float(b[tuple(i)]) for i in idx[:3]]}")
    if k == 3 and mode == 0:
        # the planner is free to feed a 3x3 differently in the two plans (halo patch / parity planes / one box per tap add the
        # nine taps in another order): the last fp16 bit may differ -- every forced-mode case below is compared bit for bit
        assert np.allclose(z1.astype(np.float32), y1.astype(np.float32), rtol=2e-3, atol=2e-3) or not store_first
        assert np.allclose(z2.astype(np.float32), y2.astype(np.float32), rtol=4e-3, atol=4e-3), where(z2, y2)
    else:
        if store_first:
            assert np.array_equal(z1, y1), "first conv's output differs from the two-launch plan: " + where(z1, y1)
        assert np.array_equal(z2, y2), "chained conv output differs from the two-launch plan: " + where(z2, y2)
    if not store_first:
        assert (z1 == 7.0).all(), "store_first=0 must leave the first conv's destination untouched"
    # and against fp32 torch
    xin = x.float()
    if up_c:
        xin = torch.cat((xu.float().repeat_interleave(2, 1).repeat_interleave(2, 2), xin), -1)
    r1 = F.silu(F.conv2d(xin.permute(0, 3, 1, 2), w1, b1, stride, k // 2)).half().float()
    r2 = F.silu(F.conv2d(r1[:, :c_in2], w2, b2)).permute(0, 2, 3, 1)
    got = torch.from_numpy(z2).float()[..., dst2_off:dst2_off + c2]
    err = (got - r2).abs()
    assert not (err > 2e-2 + 1e-2 * r2.abs()).any(), f"max err {err.max():.4f}"
    if dst2_ct != c2:
        mask = torch.ones(dst2_ct, dtype=torch.bool)
        mask[dst2_off:dst2_off + c2] = False
        assert (torch.from_numpy(z2).float()[..., mask] == 5.0).all(), "chained conv wrote outside its channel slice"


# (B, H, W, c_in, c1, c_in2, c2): the C3 cv1||cv2 -> m.0.cv1 pairs of YOLOv5m and a few ragged shapes
CASES = [
    (2, 40, 40, 96, 96, 48, 48),      # C3 of layer 2 (160^2 in the model)
    (2, 40, 40, 192, 192, 96, 96),    # C3 of layer 4: 384 accumulator columns + one chained stage
    (3, 20, 28, 64, 64, 64, 64),      # whole output consumed, ragged M (not a multiple of 128)
    (1, 16, 16, 48, 96, 32, 16),      # small K / N of the chained conv
    (5, 24, 24, 96, 128, 128, 256),   # two K groups, 256 chained output channels
    (4, 80, 80, 96, 96, 48, 48),      # many tiles per CTA (persistent loop, staging ring wrap-around)
]


@pytest.mark.parametrize("B,H,W,ci,c1,ci2,c2", CASES)
def test_chain_equals_two_launches(ctx, B, H, W, ci, c1, ci2, c2):
    run_chain_case(ctx, B, H, W, ci, c1, ci2, c2)


@pytest.mark.parametrize("mblk", [1, 2])
def test_chain_mblk_variants(ctx, mblk):
    run_chain_case(ctx, 3, 40, 40, 96, 96, 48, 48, mblk=mblk, seed=3)


def test_chain_without_storing_the_first_output(ctx):
    run_chain_case(ctx, 2, 40, 40, 96, 96, 96, 48, store_first=False, seed=4)


def test_chain_into_a_channel_slice(ctx):
    run_chain_case(ctx, 2, 20, 20, 96, 96, 48, 48, dst2_ct=96, dst2_off=48, seed=5)


def test_chain_behind_a_folded_upsample_conv(ctx):
    """C3 of layer 17: cv1||cv2 reads [upsample2x(low-res) | skip] through the stride-0 TMA map (TAP feed), m.0.cv1 chained."""
    run_chain_case(ctx, 2, 40, 40, 384, 192, 96, 96, up_c=192, seed=6)


# 3x3 producers: halo patches (s1), parity planes (s2), one box per tap; blocks that straddle images in the stack of images
@pytest.mark.parametrize("B,H,W,ci,c1,ci2,c2,stride,mode,mblk", [
    (3, 40, 40, 48, 96, 96, 96, 2, 0, 0),     # layer 1 -> C3 of layer 2's cv1||cv2 (320^2 -> 160^2 in the model)
    (2, 40, 40, 96, 192, 192, 192, 2, 0, 0),  # layer 3 -> C3 of layer 4's cv1||cv2: one stage each (feed chosen by the planner)
    (2, 40, 40, 96, 192, 192, 192, 2, 1, 1),  # the same with the parity-plane feed forced: bit for bit
    (5, 20, 20, 96, 96, 96, 48, 1, 1, 1),     # 3x3/s1 halo producer, maps shorter than a block pair
    (3, 24, 40, 64, 64, 32, 80, 1, 1, 2),
    (2, 40, 40, 48, 96, 96, 96, 2, 1, 2),     # forced HALO2, two M blocks per tile
    (2, 32, 32, 64, 64, 64, 64, 1, 2, 1),     # one box per tap
])
def test_chain_behind_3x3_producers(ctx, B, H, W, ci, c1, ci2, c2, stride, mode, mblk):
    run_chain_case(ctx, B, H, W, ci, c1, ci2, c2, k=3, stride=stride, mode=mode, mblk=mblk, seed=7)


def test_chain_3x3_s2_without_storing_the_first_output(ctx):
    """Layer 1 -> cv1||cv2 of the first C3 as the model uses it: x1 never exists in memory."""
    run_chain_case(ctx, 4, 80, 80, 48, 96, 96, 96, k=3, stride=2, store_first=False, seed=8)


def test_chain_refuses_what_it_cannot_host(ctx):
    ctx.plan_begin(2, 64, 64, 3)
    a, b, c2 = ctx.buffer(20, 20, 384), ctx.buffer(20, 20, 384), ctx.buffer(20, 20, 192)
    with pytest.raises(engine.AquaError):  # 384 output channels: the producer's tile does not hold every channel
        ctx.conv_chain((a, 0), (b, 0), torch.zeros(384, 384, 1, 1), torch.zeros(384), 1, 1, 2, (c2, 0), torch.zeros(192, 192, 1, 1), torch.zeros(192), 2)
    with pytest.raises(engine.AquaError):  # the chained conv must start at the first channel of the producer's slice
        d1 = engine.ConvDesc(a, 0, 96, b, 0, 96, 1, 1, 2, -1, 0, 1, 1, 0, 0, 0, 0, -1, 0, 0)
        d2 = engine.ConvDesc(b, 16, 48, c2, 0, 48, 1, 1, 2, -1, 0, 1, 1, 0, 0, 0, 0, -1, 0, 0)
        ctx.call("aq_plan_conv_chain", engine.C.byref(d1), engine.C.byref(d2), 1)


def test_model_with_chained_c3_equals_plain_plan(monkeypatch):
    """AQUA_CHAIN=1: the whole detector with the eligible C3 blocks chained gives bit-identical predictions and detections."""
    from aquaculture_b200 import tiles, weights
    sd = weights.synthetic_state(3, 0)
    imgs = torch.from_numpy(tiles.synthetic_tiles(range(8), 640, 640, seed=1)).pin_memory()
    out = {}
    for flag in ("0", "1", "c3"):
        monkeypatch.setenv("AQUA_CHAIN", flag)
        det = engine.TileDetector(sd, 3, batch=8, imgsz=(640, 640))
        try:
            res = det.detect(imgs, bgr=True)
            out[flag] = (det.ctx.pred(), res.dets.copy(), res.counts.copy(), det.n_launches, list(det.chained))
        finally:
            det.close()
    for flag in ("1", "c3"):
        assert out[flag][4], "nothing was chained"
        assert out[flag][3] == out["0"][3] - len(out[flag][4])
        assert np.array_equal(out["0"][0], out[flag][0]) and np.array_equal(out["0"][2], out[flag][2]) and np.array_equal(out["0"][1], out[flag][1])
    assert any("model.1 ->" in c for c in out["1"][4]) and all("->" in c and "m.0.cv1" in c for c in out["c3"][4])
