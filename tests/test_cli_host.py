"""CPU: host-side logic of the drop-in CLI (flag surface, output dir rules, letterbox shape, sharding)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from aquaculture_b200 import detect_cli as cli
from aquaculture_b200 import tiles
from oracle import letterbox as LB

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_command_line_parses():
    # the exact flags the reference README uses [REF README.md:77]
    opt = cli.parse_opt(["--weights", "output/model_weights/multilabel_farms_exp2.pt", "--source", "data/jpegs",
                         "--nosave", "--save-txt", "--save-conf"])
    assert opt.imgsz == [640, 640] and opt.conf_thres == 0.25 and opt.iou_thres == 0.45 and opt.max_det == 1000
    assert opt.nosave and opt.save_txt and opt.save_conf and opt.project == "runs/detect" and opt.name == "exp"
    cli.check_unsupported(opt)


def test_all_upstream_flags_accepted():
    opt = cli.parse_opt(["--weights", "a.pt", "--source", "x", "--data", "d.yaml", "--img", "1280", "--conf-thres", "0.3",
                         "--iou-thres", "0.5", "--max-det", "300", "--device", "0", "--save-txt", "--save-conf", "--nosave",
                         "--classes", "0", "2", "--agnostic-nms", "--project", "p", "--name", "n", "--exist-ok",
                         "--line-thickness", "2", "--hide-labels", "--hide-conf", "--half", "--vid-stride", "2"])
    assert opt.imgsz == [1280, 1280] and opt.classes == [0, 2] and opt.agnostic_nms
    assert cli.parse_opt(["--img-size", "640", "480"]).imgsz == [640, 480]


@pytest.mark.parametrize("flag", ["--augment", "--visualize", "--save-crop", "--dnn", "--update"])
def test_unsupported_modes_error_explicitly(flag):
    opt = cli.parse_opt(["--weights", "a.pt", "--nosave", "--save-txt", flag])
    with pytest.raises(SystemExit, match="unsupported"):
        cli.check_unsupported(opt)


def test_stream_sources_and_cpu_device_rejected():
    with pytest.raises(SystemExit, match="only image"):
        cli.check_unsupported(cli.parse_opt(["--weights", "a.pt", "--nosave", "--save-txt", "--source", "rtsp://x"]))
    opt = cli.parse_opt(["--weights", "synthetic", "--nosave", "--save-txt", "--device", "cpu", "--synthetic", "2"])
    with pytest.raises(SystemExit, match="no CPU fallback"):
        cli.run(opt, [])


def test_check_img_size_rounds_up():
    assert cli.check_img_size([640, 640]) == [640, 640]
    assert cli.check_img_size([1000, 650]) == [1024, 672]


def test_increment_path(tmp_path):
    p = tmp_path / "exp"
    assert cli.increment_path(p) == p
    p.mkdir()
    assert cli.increment_path(p).name == "exp2"
    (tmp_path / "exp2").mkdir()
    assert cli.increment_path(p).name == "exp3"
    assert cli.increment_path(p, exist_ok=True) == p


def test_list_images_sorted_and_filtered(tmp_path):
    for n in ["b_1_0_0.jpeg", "a_1_0_0.jpeg", "notes.txt", "c_1_0_0.PNG"]:
        (tmp_path / n).write_bytes(b"x")
    got = [os.path.basename(f) for f in cli.list_images(str(tmp_path))]
    assert got == ["a_1_0_0.jpeg", "b_1_0_0.jpeg", "c_1_0_0.PNG"]
    with pytest.raises(FileNotFoundError):
        cli.list_images(str(tmp_path / "missing"))


@pytest.mark.parametrize("hw", [(1024, 1024), (640, 640), (600, 1000), (333, 517), (2000, 900)])
def test_net_shape_matches_oracle_letterbox(hw):
    im = np.zeros((*hw, 3), np.uint8)
    assert cli.net_shape_for(hw[0], hw[1], (640, 640)) == LB.letterbox(im, (640, 640), auto=True).shape[:2]


def test_sharding_covers_each_tile_once():
    for n, w in [(10, 1), (10, 4), (200000, 8), (7, 8)]:
        seen = sorted(i for r in range(w) for i in tiles.shard_indices(n, r, w))
        assert seen == list(range(n))
    stem = tiles.tile_stem(12345)
    f = stem.split("_")
    assert len(f) == 4 and f[0][-4:] == "2020"      # what geocode_results.py:89,143 parses


def test_synthetic_tiles_are_reproducible():
    a = tiles.synthetic_tile(5, 64, 64, 0)
    assert np.array_equal(a, tiles.synthetic_tiles([5], 64, 64, 0)[0]) and not np.array_equal(a, tiles.synthetic_tile(6, 64, 64, 0))


def test_counter_gather_gloo_world2(tmp_path):
    """The path's only collective (per-rank counters) on 2 ranks over gloo."""
    script = tmp_path / "g.py"
    script.write_text(
        "import os, torch, torch.distributed as dist\n"
        "import sys; sys.path.insert(0, %r)\n"
        "from aquaculture_b200 import tiles\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "mine = tiles.shard_indices(101, r, w)\n"
        "t = torch.tensor([len(mine), sum(mine), r], dtype=torch.int64)\n"
        "allc = [torch.zeros_like(t) for _ in range(w)]\n"
        "dist.all_gather(allc, t)\n"
        "tot = torch.stack(allc).sum(0).tolist()\n"
        "assert tot[0] == 101 and tot[1] == sum(range(101)), tot\n"
        "dist.destroy_process_group()\n" % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                       capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stderr[-2000:]


def test_ranks_exit_code_reports_signals_and_failures(capsys):
    """A rank killed by a signal (negative returncode) must not hide behind the other ranks' zeros."""
    assert cli.ranks_exit_code([0, 0, 0, 0]) == 0
    assert cli.ranks_exit_code([0, -11, 0, 0]) == 139
    assert cli.ranks_exit_code([0, 0, -9]) == 137
    assert cli.ranks_exit_code([0, 3, 0]) == 3
    assert cli.ranks_exit_code([2, -11]) == 2
    err = capsys.readouterr().err
    assert "rank 1 was killed by signal 11" in err and "rank 1 exited with status 3" in err


def test_run_dir_resolved_once_under_a_launcher_gloo_world2(tmp_path):
    """Under torchrun rank 0 resolves <project>/<name>N once and broadcasts it: both ranks get the same FRESH directory
    (an existing exp is never appended to), on 2 ranks over gloo."""
    (tmp_path / "runs" / "exp").mkdir(parents=True)
    script = tmp_path / "r.py"
    script.write_text(
        "import os, sys, torch.distributed as dist\n"
        "sys.path.insert(0, %r)\n"
        "from aquaculture_b200 import detect_cli as cli\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "opt = cli.parse_opt(['--weights', 'synthetic', '--nosave', '--save-txt', '--project', %r, '--name', 'exp'])\n"
        "d = cli.resolve_run_dir(opt, r, w, dist)\n"
        "box = [None, None]\n"
        "dist.all_gather_object(box, str(d))\n"
        "assert box[0] == box[1] and box[0].endswith('exp2') and os.path.isdir(box[0]), box\n"
        "dist.destroy_process_group()\n" % (ROOT, str(tmp_path / "runs")))
    env = {k: v for k, v in os.environ.items() if k != "AQ_SPAWNED"}
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29733", str(script)],
                       capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stderr[-2000:]


def test_load_checkpoint_reads_a_whole_module_yolov5_pickle(tmp_path):
    """Upstream saves the nn.Module itself ({'model': DetectionModel.half(), 'ema': None, ...}, classes under
    models.yolo / models.common).  weights.load_checkpoint must read nc, names, anchors and every conv / BN tensor from
    such a file WITHOUT yolov5 on sys.path (it installs shim classes for the unpickler).  The file is built from the
    oracle's module classes re-homed to upstream's module paths."""
    import sys
    import types
    import torch
    from aquaculture_b200 import weights
    from oracle import yolov5m as O
    sd0 = weights.synthetic_state(3, 4)
    model = O.build(sd0, 3, fuse=False)
    homes = {O.Conv: "models.common", O.C3: "models.common", O.Bottleneck: "models.common", O.SPPF: "models.common",
             O.Concat: "models.common", O.Detect: "models.yolo", O.YOLOv5m: "models.yolo"}
    fake = {name: types.ModuleType(name) for name in ("models", "models.common", "models.yolo")}
    rehomed = {}
    for cls, home in homes.items():
        name = "DetectionModel" if cls is O.YOLOv5m else cls.__name__
        rehomed[cls] = type(name, (cls,), {"__module__": home, "__qualname__": name})
        setattr(fake[home], name, rehomed[cls])
    for m in model.modules():
        if type(m) in rehomed:
            m.__class__ = rehomed[type(m)]
    model.names = {0: "circle_farm", 1: "square_farm", 2: "other_farm"}
    path = tmp_path / "multilabel_farms_exp2.pt"
    sys.modules.update(fake)
    try:
        torch.save({"epoch": -1, "best_fitness": None, "model": model.half(), "ema": None, "updates": None, "optimizer": None,
                    "opt": {"imgsz": 640}, "date": "2023-01-01"}, path)
    finally:
        for name in fake:
            sys.modules.pop(name, None)
    assert "models" not in sys.modules and "models.yolo" not in sys.modules
    sd, nc, names = weights.load_checkpoint(str(path))
    assert nc == 3 and names == ["circle_farm", "square_farm", "other_farm"]
    assert "models" not in sys.modules  # the shims do not leak
    for k, v in sd0.items():
        if "num_batches_tracked" in k:
            continue
        assert k in sd and sd[k].dtype == torch.float32, k
        assert torch.equal(sd[k], v.half().float()), k  # .half() on save, .float() on load: upstream's round trip
    # BN folding of the loaded state == the oracle's independent fuse() of the same fp16-rounded parameters
    fw = weights.fold_bn(sd, nc)
    ref = O.build(sd, 3, fuse=True)
    w, b = fw["model.2.m.0.cv2"]
    assert torch.allclose(w, ref.model[2].m[0].cv2.conv.weight, atol=1e-6) and torch.allclose(b, ref.model[2].m[0].cv2.conv.bias, atol=1e-6)
    # a pickle of another architecture is refused, not guessed at
    torch.save({"model": torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3))}, tmp_path / "other.pt")
    with pytest.raises((ValueError, KeyError)):
        weights.load_checkpoint(str(tmp_path / "other.pt"))


def test_raster_tile_flags():
    """Additive flags for SURVEY 8f-3 / 8f-5: --raster-tile T (windows cut on the GPU), --skip-blank (needs the former)."""
    opt = cli.parse_opt(["--weights", "a.pt", "--nosave", "--save-txt", "--raster-tile", "1024", "--skip-blank"])
    assert opt.raster_tile == 1024 and opt.skip_blank
    cli.check_unsupported(opt)
    with pytest.raises(SystemExit, match="needs --raster-tile"):
        cli.check_unsupported(cli.parse_opt(["--weights", "a.pt", "--nosave", "--save-txt", "--skip-blank"]))
    with pytest.raises(SystemExit, match="--raster-tile"):
        cli.check_unsupported(cli.parse_opt(["--weights", "synthetic", "--nosave", "--save-txt", "--raster-tile", "64", "--synthetic", "4"]))


def test_jpeg_decode_modes_parse():
    base = ["--weights", "w.pt", "--source", "data/jpegs", "--nosave", "--save-txt", "--save-conf"]
    assert cli.parse_opt(base).jpeg_decode == "device"           # default: the whole decoder on the GPU
    for mode in ("device", "hybrid", "host"):
        assert cli.parse_opt(base + ["--jpeg-decode", mode]).jpeg_decode == mode
    with pytest.raises(SystemExit):
        cli.parse_opt(base + ["--jpeg-decode", "cpu"])
