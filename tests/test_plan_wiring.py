"""Host-side graph wiring vs the oracle model (CPU): the plan TileDetector builds, executed by the torch
plan emulator in fp32, must reproduce oracle/yolov5m.py's decoded predictions."""
import torch

from aquaculture_b200 import engine
from oracle import yolov5m
from plan_emulator import PlanEmulator, space_to_depth_input


import pytest


@pytest.mark.parametrize("chain", ["0", "1"])
def test_plan_matches_oracle(synthetic_sd, chain, monkeypatch):
    """chain = 1: the eligible C3 blocks are wired with the first Bottleneck's 1x1 chained behind cv1 || cv2 (AQUA_CHAIN)."""
    monkeypatch.setenv("AQUA_CHAIN", chain)
    torch.manual_seed(3)
    x = torch.rand(2, 3, 128, 160)
    emu = PlanEmulator()
    det = engine.TileDetector(synthetic_sd, 3, batch=2, imgsz=(128, 160), ctx=emu)
    assert bool(det.chained) == (chain == "1")
    pred = emu.run(space_to_depth_input(x))
    with torch.no_grad():
        ref = yolov5m.build(synthetic_sd, 3)(x)
    assert pred.shape == ref.shape == (2, det.rows, 8)
    # weights are fp16-rounded in the plan, fp32 in the oracle: small relative differences only
    err = (pred - ref).abs()
    scale = ref.abs().clamp(min=1.0)
    assert (err / scale).max() < 5e-2, (err / scale).max()  # wiring bugs give O(1) errors
    assert (err[..., 4:] ).max() < 5e-3


def test_stem_space_to_depth_exact():
    torch.manual_seed(0)
    w = torch.randn(8, 3, 6, 6)
    x = torch.rand(1, 3, 32, 48)
    ref = torch.nn.functional.conv2d(x, w, None, 2, 2)
    xs = space_to_depth_input(x).permute(0, 3, 1, 2)
    out = torch.nn.functional.conv2d(xs, engine.pack_stem_s2d(w), None, 1, 1)
    assert torch.allclose(out, ref, atol=1e-5)
