"""CPU: the oracle pinned against golden vectors from the genuine third-party kernels (tests/golden, made
by tools/make_golden.py) and against those kernels live (torchvision.ops.nms, cv2)."""
import os

import numpy as np
import pytest
import torch

from aquaculture_b200 import arch
from oracle import labels as OL
from oracle import letterbox as LB
from oracle import nms as ONMS
from oracle import yolov5m

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_architecture_selfcheck():
    # upstream README: YOLOv5m 21.2 M params, 49.0 GFLOPs @640 (nc=80)
    m = yolov5m.YOLOv5m(80)
    n = sum(p.numel() for p in m.parameters())
    assert abs(n / 1e6 - 21.19) < 0.02
    assert len(arch.iter_convs(3)) == 82
    assert abs(arch.conv_flops_per_tile(640, 640, 3) / 1e9 - 47.879) < 0.01
    assert abs(arch.conv_flops_per_tile(1280, 1280, 3) / 1e9 - 191.516) < 0.01
    assert abs(arch.conv_flops_per_tile(640, 640, 80) / 1e9 - 48.872) < 0.01
    assert arch.detect_rows(640, 640) == 25200 and arch.detect_rows(1280, 1280) == 100800


def test_nms_core_against_golden_torchvision():
    z = np.load(os.path.join(GOLD, "nms_torchvision.npz"))
    names = sorted({k.rsplit("_", 1)[0] for k in z.files})
    assert len(names) == 7
    for n in names:
        thr = float(n.split("_")[-1])
        keep = ONMS.greedy_nms_numpy(z[n + "_boxes"], z[n + "_scores"], thr)
        assert np.array_equal(keep, z[n + "_keep"]), n


def test_nms_core_live_torchvision():
    import torchvision
    g = torch.Generator().manual_seed(0)
    for n, q in [(1, None), (2, None), (700, None), (400, 8)]:
        xy = torch.rand(n, 2, generator=g) * 300
        wh = torch.rand(n, 2, generator=g) * 90 + 1
        boxes = torch.cat((xy - wh / 2, xy + wh / 2), 1)
        scores = torch.rand(n, generator=g)
        if q:
            scores = (scores * q).round() / q
        for thr in (0.45, 0.1, 0.7):
            assert np.array_equal(ONMS.greedy_nms_numpy(boxes.numpy(), scores.numpy(), thr),
                                  torchvision.ops.nms(boxes, scores, thr).numpy())


def test_nms_driver_semantics():
    # obj filter, conf = obj*cls, first-max class, per-class offsets, max_det, empty image
    pred = torch.zeros(2, 6, 7)
    pred[0, 0] = torch.tensor([50, 50, 20, 20, 0.9, 0.8, 0.1])
    pred[0, 1] = torch.tensor([51, 50, 20, 20, 0.9, 0.7, 0.1])   # same class, overlaps -> suppressed
    pred[0, 2] = torch.tensor([51, 50, 20, 20, 0.9, 0.1, 0.6])   # other class -> kept
    pred[0, 3] = torch.tensor([200, 200, 20, 20, 0.2, 0.9, 0.9])  # obj below threshold
    pred[0, 4] = torch.tensor([300, 300, 20, 20, 0.6, 0.4, 0.4])  # conf 0.24 below threshold
    pred[0, 5] = torch.tensor([400, 400, 20, 20, 0.9, 0.5, 0.5])  # class tie -> class 0
    out, rows = ONMS.non_max_suppression(pred, return_rows=True)
    assert rows[0].tolist() == [0, 2, 5] and out[0][:, 5].tolist() == [0.0, 1.0, 0.0]
    assert out[1].shape == (0, 6)
    assert torch.allclose(out[0][0, :4], torch.tensor([40.0, 40, 60, 60]))
    assert len(ONMS.non_max_suppression(pred, max_det=2)[0]) == 2
    assert ONMS.non_max_suppression(pred, agnostic=True, return_rows=True)[1][0].tolist() == [0, 5]
    assert ONMS.non_max_suppression(pred, classes=[1], return_rows=True)[1][0].tolist() == [2]


def test_resize_against_golden_cv2():
    z = np.load(os.path.join(GOLD, "resize_cv2.npz"))
    for n in sorted({k.rsplit("_", 1)[0] for k in z.files}):
        dst = z[n + "_dst"]
        got = LB.resize_bilinear_u8(z[n + "_src"], dst.shape[1], dst.shape[0])
        assert np.array_equal(got, dst), f"{n}: {np.count_nonzero(got != dst)} samples differ"


@pytest.mark.parametrize("hw,new", [((1024, 1024), 640), ((600, 1000), 640), ((300, 500), 640), ((640, 640), 640),
                                    ((50, 37), 640)])
def test_letterbox_live_cv2(hw, new):
    rng = np.random.default_rng(3)
    im = rng.integers(0, 256, (*hw, 3), dtype=np.uint8)
    for auto in (True, False):
        a = LB.letterbox(im, (new, new), auto=auto, use_cv2=True)
        b = LB.letterbox(im, (new, new), auto=auto, use_cv2=False)
        assert np.array_equal(a, b)
        assert a.shape[0] % 32 == 0 and a.shape[1] % 32 == 0
        if not auto:
            assert a.shape[:2] == (new, new)


def test_letterbox_reference_tile_geometry():
    # 1024^2 tiles [REF src/utils.py:18-19] -> r = 0.625, 640^2, no padding
    assert LB.letterbox_geometry(1024, 1024, (640, 640)) == (640, 640, 0, 0, 0, 0)
    assert LB.letterbox_geometry(640, 640, (640, 640)) == (640, 640, 0, 0, 0, 0)
    nw, nh, t, b, l, r = LB.letterbox_geometry(600, 1000, (640, 640), auto=True)
    assert (nw, nh) == (640, 384) and (t, b, l, r) == (0, 0, 0, 0)
    assert LB.letterbox_geometry(600, 1000, (640, 640), auto=False)[2:] == (128, 128, 0, 0)


def test_label_grammar_and_consumer(tmp_path):
    det = torch.tensor([[10.2, 20.7, 30.4, 41.0, 0.873421, 0.0], [100.0, 100.0, 140.6, 120.5, 0.5, 1.0]])
    lines = OL.label_lines(det, (640, 640), (1024, 1024))
    # ascending confidence, %g, six columns
    assert lines[0].split()[0] == "1" and lines[1].split()[0] == "0"
    assert lines[1] == "0 0.0317383 0.0483398 0.0322266 0.0322266 0.873421"
    assert OL.write_labels(str(tmp_path), "ORTHO2020_12_0_5120", lines)
    assert not OL.write_labels(str(tmp_path), "ORTHO2020_12_0_6144", [])
    assert not (tmp_path / "ORTHO2020_12_0_6144.txt").exists()   # zero detections -> NO file
    # the consumer's own parsing [REF src/process_yolo/geocode_results.py:140-172]
    path = str(tmp_path / "ORTHO2020_12_0_5120.txt")
    inf_bboxs = np.loadtxt(path)
    assert inf_bboxs.shape == (2, 6)
    _, bbox_ind, x_offset, y_offset = os.path.basename(path).replace(".txt", "").split("_")
    year = int(os.path.basename(path).replace(".txt", ".jpeg").split("_")[0][-4:])
    assert (int(bbox_ind), int(x_offset), int(y_offset), year) == (12, 0, 5120, 2020)
    REVERSE_CLASS_MAPPING = {0: "circle_farm", 1: "square_farm", 2: "triangle_farm", 3: "other_farm", 4: "rectangle_farm"}
    for bbox in inf_bboxs:
        xmin = int(1024 * (bbox[1] - bbox[3] / 2))
        assert 0 <= xmin <= 1024 and 0 < bbox[5] <= 1
        assert REVERSE_CLASS_MAPPING[bbox[0]]
    one = np.loadtxt(_one_line(tmp_path))
    assert one.ndim == 1 and one.shape == (6,)


def _one_line(tmp_path):
    p = tmp_path / "single.txt"
    p.write_text("0 0.5 0.25 0.0195312 0.0205078 0.873421\n")
    return str(p)


def test_scale_boxes_clip_round():
    det = torch.tensor([[-5.0, 3.3, 700.0, 650.0, 0.9, 2.0]])
    lines = OL.label_lines(det, (640, 640), (1024, 1024))
    c, x, y, w, h, conf = (float(v) for v in lines[0].split())
    assert (c, conf) == (2.0, 0.9) and x == 0.5 and w == 1.0 and abs(y - (5 + 1024) / 2 / 1024) < 1e-6


def test_u8_over_255_fp16_equals_reciprocal_multiply():
    """preprocess_kernel evaluates u/255 -> fp16 as half(float(u) * (1/255f)); upstream computes `.half() / 255`
    (detect.py::run).  Exhaustive over the 256 inputs: all three roundings agree."""
    u = np.arange(256, dtype=np.float32)
    by_mul = (u * np.float32(1.0 / 255.0)).astype(np.float16)
    by_div = (u / np.float32(255.0)).astype(np.float16)
    upstream = (torch.arange(256, dtype=torch.uint8).half() / 255).numpy()
    assert (by_mul == by_div).all() and (by_mul == upstream).all()


def test_blank_oracle_matches_reference_source():
    """oracle/blank.py vs outputs of the reference's own is_blank / is_partly_blank source on seeded tiles
    (tests/golden/blank_ref.npz, made by tools/make_golden_blank.py from /root/reference/src/utils.py:325-369)."""
    from oracle import blank as OB
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "blank_ref.npz"))
    assert g["is_blank"].any() and (~g["is_blank"]).any() and g["is_partly_blank"].any() and (~g["is_partly_blank"]).any()
    for i, t in enumerate(g["tiles"]):
        s = OB.tile_stats(t)
        assert (s[0], s[1]) == tuple(g["extrema"][i]), f"tile {i}: L extrema {s[:2]} vs PIL {g['extrema'][i]}"
        assert OB.is_blank(t) == bool(g["is_blank"][i]), f"tile {i}: is_blank"
        assert OB.is_partly_blank(t) == bool(g["is_partly_blank"][i]), f"tile {i}: is_partly_blank"


def test_window_grid_order_matches_reference_tiler():
    """x outer, y inner, names <x>_<y> [REF src/load_data/tile_tifs.py:33-41]."""
    from oracle import blank as OB
    from aquaculture_b200 import engine
    assert OB.windows(3072, 2048, 1024) == [(0, 0), (0, 1024), (0, 2048), (1024, 0), (1024, 1024), (1024, 2048)]
    assert engine.window_grid(3072, 2048, 1024) == OB.windows(3072, 2048, 1024)
    assert engine.window_grid(2500, 2100, 1024) == [(0, 0), (0, 1024), (1024, 0), (1024, 1024)]  # ragged strips left out


def test_parity_compare_image_proves_flips_and_flags_real_differences():
    """oracle/parity.py::compare_image (the checker behind tests/test_parity_b64.py and bench.py's parity leg): predictions that
    differ by fp16-sized noise yield only proven threshold flips; a detection that is genuinely wrong, a box that moved, or
    an NMS that did something else is reported."""
    from oracle import nms as N, parity
    flips = 0
    for seed in range(4):
        g = torch.Generator().manual_seed(seed)
        rows = 3000
        pred = torch.zeros(rows, 8)
        pred[:, 0:2] = torch.rand(rows, 2, generator=g) * 640
        pred[:, 2:4] = torch.rand(rows, 2, generator=g) * 60 + 10
        pred[:, 4] = torch.rand(rows, generator=g) * 0.5
        pred[:, 5:] = torch.rand(rows, 3, generator=g)
        noisy = pred + torch.randn(rows, 8, generator=g) * 1e-3 * torch.tensor([0.5, 0.5, 0.5, 0.5, 1, 1, 1, 1])
        (d,), _ = N.non_max_suppression(noisy[None], return_rows=True)
        m = parity.compare_image(noisy.numpy(), d.numpy(), pred.numpy())
        assert m["nms_exact"] and m["unexplained"] == 0 and m["box_err"] < 1e-3 and m["conf_err"] < 5e-3, m
        flips += m["flips_conf"] + m["flips_iou"] + m["flips_order"] + m["flips_cascade"] + m["flips_class"]
        # (1) a detection dropped by the path under test although it is far from every threshold
        k = int(torch.argmax(d[:, 4]))
        m1 = parity.compare_image(noisy.numpy(), torch.cat((d[:k], d[k + 1:])).numpy(), pred.numpy())
        assert not m1["nms_exact"]
        # (2) a prediction row moved by three pixels: caught by the row-matched box error
        moved = noisy.clone()
        row = int(torch.argmax(noisy[:, 4] * noisy[:, 5:].amax(1)))
        moved[row, 0] += 3.0
        (d2,), _ = N.non_max_suppression(moved[None], return_rows=True)
        m2 = parity.compare_image(moved.numpy(), d2.numpy(), pred.numpy())
        assert m2["box_err"] > 1e-2 or m2["unexplained"] > 0, m2
        # (3) a confident row whose objectness collapsed: one-sided and NOT at a threshold
        gone = noisy.clone()
        gone[row, 4] = 0.01
        (d3,), _ = N.non_max_suppression(gone[None], return_rows=True)
        m3 = parity.compare_image(gone.numpy(), d3.numpy(), pred.numpy())
        assert m3["unexplained"] >= 1 and m3["conf_err"] > 5e-3, m3
    assert flips > 0, "the noise never produced a flip: the proof obligations were not exercised"
