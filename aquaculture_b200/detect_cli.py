"""Drop-in command line for the reference's inference step

    python3 yolov5/detect.py --weights ...exp2.pt --source data/jpegs --nosave --save-txt --save-conf
                                                                            [REF README.md:77]

Mirrors upstream `detect.py::{parse_opt, main, run}` [UPSTREAM-RECALLED, SURVEY.md §8 a-1 / §8b]: same flag
names and defaults, same output layout `<project>/<name>{,2,3..}/labels/<stem>.txt`, same label grammar
(consumer: [REF src/process_yolo/geocode_results.py:140-172]).  What differs is underneath: tiles are batched,
every arithmetic step runs in libaqua_b200.so on a B200, and `--device 0,1,...` shards the sorted tile list
over GPUs (one process per GPU, no collective on the data path; NCCL gathers three counters at the end).
Modes the B200 path does not implement raise instead of falling back to another executor.
"""
from __future__ import annotations

import argparse
import glob
import os
import re
import subprocess
import sys
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import List, Optional, Tuple

IMG_FORMATS = ("bmp", "dng", "jpeg", "jpg", "mpo", "png", "tif", "tiff", "webp", "pfm")
STRIDE = 32


def parse_opt(argv=None):
    p = argparse.ArgumentParser(prog="detect.py")
    p.add_argument("--weights", nargs="+", type=str, default=["yolov5s.pt"], help="model path; 'synthetic' = seeded stand-in weights")
    p.add_argument("--source", type=str, default="data/images", help="file/dir/glob of images")
    p.add_argument("--data", type=str, default="data/coco128.yaml", help="(accepted, unused: class names come from the checkpoint)")
    p.add_argument("--imgsz", "--img", "--img-size", nargs="+", type=int, default=[640], help="inference size h,w")
    p.add_argument("--conf-thres", type=float, default=0.25)
    p.add_argument("--iou-thres", type=float, default=0.45)
    p.add_argument("--max-det", type=int, default=1000)
    p.add_argument("--device", default="", help="cuda device(s), i.e. 0 or 0,1,2,3 (tiles are sharded across them)")
    p.add_argument("--view-img", action="store_true")
    p.add_argument("--save-txt", action="store_true")
    p.add_argument("--save-conf", action="store_true")
    p.add_argument("--save-crop", action="store_true")
    p.add_argument("--nosave", action="store_true")
    p.add_argument("--classes", nargs="+", type=int)
    p.add_argument("--agnostic-nms", action="store_true")
    p.add_argument("--augment", action="store_true")
    p.add_argument("--visualize", action="store_true")
    p.add_argument("--update", action="store_true")
    p.add_argument("--project", default="runs/detect")
    p.add_argument("--name", default="exp")
    p.add_argument("--exist-ok", action="store_true")
    p.add_argument("--line-thickness", default=3, type=int)
    p.add_argument("--hide-labels", default=False, action="store_true")
    p.add_argument("--hide-conf", default=False, action="store_true")
    p.add_argument("--half", action="store_true", help="(accepted: the B200 path always computes fp16 in, fp32 accumulate)")
    p.add_argument("--dnn", action="store_true")
    p.add_argument("--vid-stride", type=int, default=1)
    # additive
    p.add_argument("--batch-size", type=int, default=64, help="tiles per GPU batch")
    p.add_argument("--synthetic", type=int, default=0, help="run on N seeded synthetic tiles instead of --source")
    p.add_argument("--synthetic-size", type=int, default=640)
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--nc", type=int, default=3, help="classes of the synthetic stand-in weights")
    p.add_argument("--raster-tile", type=int, default=0,
                   help="treat every source image as a large raster and detect on its TxT windows cut on the GPU "
                        "(replaces src/load_data/tile_tifs.py; labels are named <stem>_<x>_<y>.txt like its tiles)")
    p.add_argument("--skip-blank", action="store_true",
                   help="with --raster-tile: skip windows the reference's is_blank() calls blank (src/utils.py:325-350)")
    p.add_argument("--jpeg-decode", choices=("device", "hybrid", "host"), default="device",
                   help="device: baseline .jpg/.jpeg tiles are decoded entirely on the GPU (parallel Huffman decoding, IDCT, chroma "
                        "upsampling, colour), bit-identical to cv2.imread; the host only walks the markers.  hybrid: Huffman decoding "
                        "by the library's C parser on the io threads, the rest on the GPU (also used by `device` for files with "
                        "restart intervals).  Variants neither covers (progressive, 4:2:2, CMYK) and other formats are read with "
                        "cv2.imread as upstream does.  host: cv2.imread for everything")
    p.add_argument("--verbose", action="store_true", help="per-image log lines as upstream prints them")
    p.add_argument("--io-threads", type=int, default=8)
    opt = p.parse_args(argv)
    opt.imgsz *= 2 if len(opt.imgsz) == 1 else 1
    return opt


def check_unsupported(opt):
    bad = [f for f in ("augment", "visualize", "save_crop", "dnn", "update") if getattr(opt, f)]
    if bad:
        raise SystemExit("detect.py (B200 path): unsupported mode(s) " + ", ".join("--" + b.replace("_", "-") for b in bad) +
                         "; there is no fallback executor")
    if opt.view_img:
        print("WARNING: --view-img ignored (no window system)", file=sys.stderr)
    if not opt.nosave and not opt.save_txt:
        raise SystemExit("detect.py (B200 path): annotated-image output is not implemented; pass --nosave --save-txt "
                         "[--save-conf] as the reference does [README.md:77]")
    if opt.skip_blank and opt.raster_tile <= 0:
        raise SystemExit("detect.py (B200 path): --skip-blank needs --raster-tile (the measurement runs on the device-resident raster)")
    if opt.raster_tile < 0 or (opt.raster_tile > 0 and opt.synthetic > 0):
        raise SystemExit("detect.py (B200 path): --raster-tile takes a positive window size and real --source rasters")
    if len(opt.weights) != 1:
        raise SystemExit("detect.py (B200 path): model ensembles (several --weights) are not supported")
    src = opt.source
    if opt.synthetic <= 0 and (src.isnumeric() or src.lower().startswith(("rtsp://", "rtmp://", "http://", "https://", "screen")) or
                               src.endswith(".streams") or src.rsplit(".", 1)[-1].lower() in ("mp4", "avi", "mov", "mkv", "gif")):
        raise SystemExit("detect.py (B200 path): only image files/directories are supported as --source")


def check_img_size(imgsz, s=STRIDE):
    """upstream utils/general.py::check_img_size: round each side UP to a multiple of the max stride."""
    import math
    new = [max(int(math.ceil(x / s) * s), 0) for x in imgsz]
    if new != list(imgsz):
        print(f"WARNING: --img-size {imgsz} must be multiple of max stride {s}, updating to {new}", file=sys.stderr)
    return new


def increment_path(path, exist_ok=False) -> Path:
    """upstream utils/general.py::increment_path: runs/exp -> runs/exp2, runs/exp3, ..."""
    path = Path(path)
    if path.exists() and not exist_ok:
        for n in range(2, 9999):
            p = f"{path}{n}"
            if not os.path.exists(p):
                path = Path(p)
                break
    return path


def list_images(source: str) -> List[str]:
    """upstream utils/dataloaders.py::LoadImages: sorted glob / dir listing / single file, image suffixes only."""
    p = str(Path(source).resolve()) if not any(ch in source for ch in "*?[") else source
    if "*" in p or "?" in p or "[" in p:
        files = sorted(glob.glob(p, recursive=True))
    elif os.path.isdir(p):
        files = sorted(glob.glob(os.path.join(p, "*.*")))
    elif os.path.isfile(p):
        files = [p]
    else:
        raise FileNotFoundError(f"{p} does not exist")
    images = [x for x in files if x.rsplit(".", 1)[-1].lower() in IMG_FORMATS]
    if not images:
        raise FileNotFoundError(f"No images found in {p}. Supported formats are: {IMG_FORMATS}")
    return images


def net_shape_for(h0: int, w0: int, imgsz: Tuple[int, int]) -> Tuple[int, int]:
    """Letterboxed (h, w) upstream produces with auto=True (minimum rectangle, multiples of 32)."""
    r = min(imgsz[0] / h0, imgsz[1] / w0)
    nw, nh = int(round(w0 * r)), int(round(h0 * r))
    dw, dh = (imgsz[1] - nw) % STRIDE, (imgsz[0] - nh) % STRIDE
    top, bottom = int(round(dh / 2 - 0.1)), int(round(dh / 2 + 0.1))
    left, right = int(round(dw / 2 - 0.1)), int(round(dw / 2 + 0.1))
    return nh + top + bottom, nw + left + right


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def spawn_ranks(devices: List[int], argv: List[str]) -> int:
    """`--device 0,1,..`: one process per GPU; rank r takes tiles r, r+W, ... of the sorted list."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for r, d in enumerate(devices):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(len(devices)), LOCAL_RANK=str(d), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), AQ_SPAWNED="1")
        procs.append(subprocess.Popen([sys.executable, "-m", "aquaculture_b200.detect_cli"] + argv, env=env))
    return ranks_exit_code([p.wait() for p in procs])


def ranks_exit_code(rcs: List[int]) -> int:
    """Exit status of a multi-rank run: 0 only if EVERY rank exited 0.  A rank killed by a signal has a negative
    returncode (-11, -9 ...), which a plain max() would hide behind the other ranks' zeros -- its shard of label files
    would be missing and the run would still look fine; such a code is reported the shell way (128 + signal)."""
    bad = [(r, rc) for r, rc in enumerate(rcs) if rc != 0]
    for r, rc in bad:
        print(f"detect.py (B200 path): rank {r} " + (f"was killed by signal {-rc}" if rc < 0 else f"exited with status {rc}"),
              file=sys.stderr)
    if not bad:
        return 0
    rc = bad[0][1]
    return rc if rc > 0 else 128 - rc


def resolve_run_dir(opt, rank: int, world: int, dist=None) -> Path:
    """The run directory `<project>/<name>{,2,3..}` [upstream increment_path].  Under an external launcher (torchrun:
    WORLD_SIZE > 1, not self-spawned) rank 0 resolves it once and broadcasts it, so that every rank writes into the same
    fresh expN -- resolving it per rank would race, and reusing an existing directory would APPEND to the previous run's
    label files (aq_format_labels opens them with "a", as upstream does)."""
    if world == 1 or "AQ_SPAWNED" in os.environ:
        # self-spawned ranks were handed the parent's already-resolved directory with --exist-ok
        return increment_path(Path(opt.project) / opt.name, exist_ok=opt.exist_ok)
    box = [str(increment_path(Path(opt.project) / opt.name, exist_ok=opt.exist_ok)) if rank == 0 else None]
    if rank == 0:
        Path(box[0]).mkdir(parents=True, exist_ok=True)  # taken before any other launch can pick the same name
    dist.broadcast_object_list(box, src=0)
    return Path(box[0])


def run(opt, argv) -> int:
    check_unsupported(opt)
    rank, world, local = dist_env()
    devs = [int(d) for d in str(opt.device).replace("cuda:", "").split(",") if d.strip() != ""] if opt.device not in ("", "cpu") else [0]
    if opt.device == "cpu":
        raise SystemExit("detect.py (B200 path): --device cpu is not available; there is no CPU fallback")
    if len(devs) > 1 and "AQ_SPAWNED" not in os.environ and world == 1:
        # resolve the run directory once so all ranks write into the same labels/
        save_dir = increment_path(Path(opt.project) / opt.name, exist_ok=opt.exist_ok)
        (save_dir / "labels").mkdir(parents=True, exist_ok=True)
        argv = list(argv) + ["--project", str(save_dir.parent), "--name", save_dir.name, "--exist-ok"]
        return spawn_ranks(devs, argv)
    device = local if "AQ_SPAWNED" in os.environ or world > 1 else devs[0]

    import numpy as np
    import torch
    from . import engine, tiles, weights

    imgsz = tuple(check_img_size(opt.imgsz))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(device)
        dist.init_process_group(os.environ.get("AQ_DIST_BACKEND", "nccl"), device_id=torch.device("cuda", device))
    save_dir = resolve_run_dir(opt, rank, world, dist)
    label_dir = save_dir / "labels"
    if opt.save_txt:
        label_dir.mkdir(parents=True, exist_ok=True)
    else:
        save_dir.mkdir(parents=True, exist_ok=True)

    # ---- model
    w = opt.weights[0]
    if w == "synthetic":
        sd, nc, names = weights.synthetic_state(opt.nc, opt.seed), opt.nc, [f"class{i}" for i in range(opt.nc)]
    else:
        if not os.path.isfile(w):
            raise SystemExit(f"detect.py: weights file {w} not found")
        sd, nc, names = weights.load_checkpoint(w)

    # ---- tile list and this rank's shard (strided over the sorted list)
    if opt.synthetic > 0:
        idx = tiles.shard_indices(opt.synthetic, rank, world)
        items = [("synthetic", i) for i in idx]
        stems = [tiles.tile_stem(i) for i in idx]
    else:
        files = list_images(opt.source)
        files = [files[i] for i in tiles.shard_indices(len(files), rank, world)]
        items = [("file", f) for f in files]
        stems = [Path(f).stem for f in files]

    import cv2
    from . import jpeg
    cv2.setNumThreads(1)
    device_jpeg = opt.jpeg_decode in ("device", "hybrid") and opt.raster_tile <= 0
    raster_jpeg = opt.jpeg_decode == "device" and opt.raster_tile > 0  # a .jpeg raster is decoded straight into device memory
    n_host_decoded = [0]
    n_blank = [0]  # windows --skip-blank kept away from the network

    def load(item):
        kind, v = item
        if kind == "synthetic":
            return tiles.synthetic_tile(v, opt.synthetic_size, opt.synthetic_size, opt.seed)
        if raster_jpeg and v.rsplit(".", 1)[-1].lower() in ("jpg", "jpeg"):
            with open(v, "rb") as fh:
                blob = np.frombuffer(fh.read(), np.uint8)
            try:
                info = jpeg.probe(blob)
                if info.restart_interval == 0:
                    return engine.RawJpeg(blob, info)
            except (jpeg.JpegUnsupported, engine.AquaError):
                pass  # a variant without a device path, or a file the strict marker walk rejects: libjpeg decides
            n_host_decoded[0] += 1
            im = cv2.imdecode(blob, cv2.IMREAD_COLOR)
            if im is None:
                raise SystemExit(f"Image Not Found {v}")
            return im
        if device_jpeg and v.rsplit(".", 1)[-1].lower() in ("jpg", "jpeg"):
            with open(v, "rb") as fh:
                blob = np.frombuffer(fh.read(), np.uint8)
            try:
                if opt.jpeg_decode == "device":
                    info = jpeg.probe(blob)  # geometry only; the whole batch is prepared by one library call in flush()
                    if info.restart_interval == 0:
                        return engine.RawJpeg(blob, info)
                    # restart intervals re-align the stream: Huffman on the host for such a file
                return engine.entropy_decode_file(blob)  # host half: markers + Huffman (C, GIL released)
            except (jpeg.JpegUnsupported, engine.AquaError):
                # progressive / 4:2:2 / CMYK, or a file the strict marker walk rejects (stray bytes between segments, EOI
                # oddities): decoded on the host like upstream does -- libjpeg tolerates what it can -- and reported below
                n_host_decoded[0] += 1
                im = cv2.imdecode(blob, cv2.IMREAD_COLOR)
                if im is None:
                    raise SystemExit(f"Image Not Found {v}")
                return im
        im = cv2.imread(v)  # BGR, as upstream LoadImages
        if im is None:
            raise SystemExit(f"Image Not Found {v}")
        return im

    detectors = {}

    def detector_for(net_hw):
        if net_hw not in detectors:
            detectors[net_hw] = engine.TileDetector(sd, nc, batch=opt.batch_size, imgsz=net_hw, device=device,
                                                    conf_thres=opt.conf_thres, iou_thres=opt.iou_thres, max_det=opt.max_det,
                                                    agnostic=opt.agnostic_nms, classes=opt.classes)
        return detectors[net_hw]

    t_start = time.perf_counter()
    dt = [0.0, 0.0]  # load+stage, gpu (pre+inference+NMS) + label write
    seen = n_files = n_dets = 0
    pool = ThreadPoolExecutor(opt.io_threads)
    pack_pool = ThreadPoolExecutor(max(2, opt.io_threads // 4))  # staging copies: not queued behind the prefetched loads
    B = opt.batch_size
    pos = 0
    pending = {}  # (h0, w0) -> list of (image, stem)

    deferred = []  # the batch in flight: (detector, ticket, group, shape0, net_hw, scan) -- scan = (pinned per-tile decoder
                   # status, blobs) of an all-device JPEG batch, else None

    def finish():
        nonlocal seen, n_files, n_dets
        if not deferred:
            return
        det, ticket, group, shape0, net_hw, scan = deferred.pop()
        t1 = time.perf_counter()
        res = det.detect_wait(ticket)
        if scan is not None and scan[0].any():
            # the device decoder reported these scans as damaged (truncated, bad code): they go through cv2 / libjpeg, which
            # pads or warns and still returns pixels where it can -- what upstream's cv2.imread would have done
            bad = [i for i, v in enumerate(scan[0]) if v]
            redo = []
            for i in bad:
                im = cv2.imdecode(scan[1][i], cv2.IMREAD_COLOR)
                if im is None or im.shape[:2] != tuple(shape0):
                    raise SystemExit(f"Image Not Found {group[i][1]} (corrupt JPEG scan, device status {int(scan[0][i])})")
                redo.append(im)
            n_host_decoded[0] += len(bad)
            print(f"WARNING: {len(bad)} damaged JPEG scan(s) re-decoded with cv2: " + ", ".join(group[i][1] for i in bad), file=sys.stderr)
            res = engine.Detections(res.dets.copy(), res.counts.copy())
            fix = det.detect(torch.from_numpy(np.stack(redo)).pin_memory(), bgr=True)  # into the (copied) slot of this batch
            det._slot ^= 1  # the batch submitted after this one keeps its result slot; alternation continues from there
            for k, i in enumerate(bad):
                res.dets[i], res.counts[i] = fix.dets[k], fix.counts[k]
        if opt.save_txt:
            paths = [str(label_dir / (g[1] + ".txt")) for g in group]
            n_files += engine.write_labels(res, net_hw, [shape0] * len(group), paths, save_conf=opt.save_conf,
                                           threads=opt.io_threads)
        dt[1] += time.perf_counter() - t1
        seen += len(group)
        n_dets += int(res.counts.sum())
        if opt.verbose:
            for g, c in zip(group, res.counts):
                print(f"image {g[1]}: {net_hw[0]}x{net_hw[1]} {int(c)} detections")

    def flush(shape0, group):
        # two-deep pipeline: this batch is submitted (H2D on the copy stream + kernels) before the previous one is
        # waited for, so image decoding / label writing on the host overlap the GPU
        is_jpeg, is_scan = shape0[0] == "jpeg", shape0[0] == "jpegscan"
        if is_jpeg or is_scan:
            _, h0, w0, ncomp, hs = shape0
            shape0 = (h0, w0)
        h0, w0 = shape0
        net_hw = net_shape_for(h0, w0, imgsz)
        det = detector_for(net_hw)
        t0 = time.perf_counter()
        scan = None
        if is_jpeg:
            desc, ent, base, quant = det.jpeg_staging(len(group), group[0][0].info.coef_count, sum(g[0].entries.size for g in group))
            total = engine.pack_jpeg_tiles([g[0] for g in group], desc, ent, base, quant, pool=pack_pool)
            dt[0] += time.perf_counter() - t0
            ticket = det.detect_submit_jpeg(desc, ent, base, total, quant, w0, h0, ncomp, hs)
        elif is_scan:
            try:
                (sc, off, ln, tab, quant, status), total, _ = engine.prepare_jpeg_batch(det, [g[0].blob for g in group], n_threads=opt.io_threads)
            except engine.AquaError as e:
                raise SystemExit(f"detect.py (B200 path): {e} (batch starting at {group[0][1]})")
            dt[0] += time.perf_counter() - t0
            ticket = det.detect_submit_jpeg_scans(sc, off, ln, total, tab, quant, status, w0, h0, ncomp, hs)
            scan = (status, [g[0].blob for g in group])
        else:
            batch = det.pinned_input(len(group), h0, w0)
            np.stack([g[0] for g in group], out=batch.numpy())
            dt[0] += time.perf_counter() - t0
            ticket = det.detect_submit(batch, bgr=True)
        finish()
        deferred.append((det, ticket, [(None, g[1]) for g in group], shape0, net_hw, scan))

    def run_raster(im, stem):
        """Windows of one raster: full TxT windows are cut, letterboxed and measured on the device; the ragged right /
        bottom strips (the reference's tiler emits them as smaller tiles) are cropped on the host and batched by shape."""
        nonlocal seen, n_files, n_dets
        T = opt.raster_tile
        finish()  # detect_windows uses the detector's result slot 0 and synchronises
        raster = None
        if isinstance(im, engine.RawJpeg):  # decoded on the device; pixels come to the host only for ragged strips
            t0 = time.perf_counter()
            raster = engine.decode_jpeg_raster(detector_for(net_shape_for(T, T, imgsz)).ctx, im.blob)
            dt[0] += time.perf_counter() - t0
            H, W = raster.shape[:2]
            im = raster.cpu().numpy() if (H % T or W % T) else None
        else:
            H, W = im.shape[:2]
        wins = engine.window_grid(H, W, T)
        if wins:
            net_hw = net_shape_for(T, T, imgsz)
            det = detector_for(net_hw)
            t0 = time.perf_counter()
            if raster is None:
                raster = torch.from_numpy(np.ascontiguousarray(im)).to(f"cuda:{device}")
            dt[0] += time.perf_counter() - t0
            live = list(range(len(wins)))
            if opt.skip_blank:
                # measure every window first; the blank ones never reach the network (the reference does not even download
                # such images [REF src/load_data/download_french_data.py:58-60])
                t1 = time.perf_counter()
                live = []
                for k0 in range(0, len(wins), B):
                    nb = min(B, len(wins) - k0)
                    st = engine.tile_stats(det.ctx, raster, bgr=True, tile=T, win0=k0, n=nb)
                    live += [k0 + i for i in range(nb) if not engine.is_blank(st[i])]
                n_blank[0] += len(wins) - len(live)
                seen += len(wins) - len(live)
                dt[1] += time.perf_counter() - t1
            for k0 in range(0, len(live), B):
                idx = live[k0:k0 + B]
                nb = len(idx)
                t1 = time.perf_counter()
                if idx == list(range(idx[0], idx[0] + nb)):
                    res = det.detect_windows(raster, T, idx[0], nb, bgr=True)
                else:
                    res = det.detect_window_list(raster, T, idx, bgr=True)
                if opt.save_txt:
                    paths = [str(label_dir / f"{stem}_{wins[k][0]}_{wins[k][1]}.txt") for k in idx]
                    n_files += engine.write_labels(res, net_hw, [(T, T)] * nb, paths, save_conf=opt.save_conf, threads=opt.io_threads)
                dt[1] += time.perf_counter() - t1
                seen += nb
                n_dets += int(res.counts.sum())
        for x in range(0, W, T):
            for y in range(0, H, T):
                if x + T <= W and y + T <= H:
                    continue
                crop = np.ascontiguousarray(im[y:y + T, x:x + T])
                key = crop.shape[:2]
                pending.setdefault(key, []).append((crop, f"{stem}_{x}_{y}"))
                if len(pending[key]) == B:
                    flush(key, pending.pop(key))

    if opt.raster_tile > 0:
        for item, stem in zip(items, stems):
            t0 = time.perf_counter()
            im = load(item)
            dt[0] += time.perf_counter() - t0
            run_raster(im, stem)
        pos = len(items)
    # the next chunk's files are read / parsed on the io threads while this chunk is grouped, submitted and its
    # predecessor's labels are written
    futs = [pool.submit(load, it) for it in items[pos:pos + B]]
    while pos < len(items):
        nxt = [pool.submit(load, it) for it in items[pos + B:pos + 2 * B]]
        t0 = time.perf_counter()
        imgs = [f.result() for f in futs]
        futs = nxt
        dt[0] += time.perf_counter() - t0
        for im, stem in zip(imgs, stems[pos:pos + B]):
            key = im.key if isinstance(im, (engine.PackedJpeg, engine.RawJpeg)) else im.shape[:2]
            pending.setdefault(key, []).append((im, stem))
            if len(pending[key]) == B:
                flush(key, pending.pop(key))
        pos += B
    for key in list(pending):
        flush(key, pending.pop(key))
    finish()
    stage_ms, stage_tiles = [0.0, 0.0, 0.0], 0
    for d in detectors.values():
        if d.stage_tiles == 0:  # a detector that saw one batch only: its (warm-up) batch is all there is
            d.stage_ms, d.stage_tiles = d.first_stage_ms, d.first_stage_tiles
        stage_ms = [a + b for a, b in zip(stage_ms, d.stage_ms)]
        stage_tiles += d.stage_tiles
        d.close()
    wall = time.perf_counter() - t_start

    # ---- the one collective: gather per-rank counters (NCCL), rank 0 prints the summary
    totals = [seen, n_files, n_dets]
    if world > 1:
        t = torch.tensor(totals, dtype=torch.int64, device="cuda")
        allc = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allc, t)
        totals = torch.stack(allc).sum(0).tolist()
        dist.destroy_process_group()
    if rank == 0:
        per = 1e3 / max(seen, 1)
        t3 = [v / max(stage_tiles, 1) for v in stage_ms]
        # upstream's line, the three stages measured with CUDA events on the device (this rank's batches) ...
        print(f"Speed: {t3[0]:.3f}ms pre-process, {t3[1]:.3f}ms inference, {t3[2]:.3f}ms NMS per image at shape "
              f"({opt.batch_size}, 3, {imgsz[0]}, {imgsz[1]})")
        # ... and what the host sees around them
        print(f"Host: {dt[0] * per:.2f}ms load, {dt[1] * per:.2f}ms wait+labels per image; {totals[0] / wall:.1f} tiles/s over {world} GPU(s)")
        if n_blank[0]:
            print(f"{n_blank[0]} blank window(s) skipped before inference (--skip-blank; this rank)")
        if n_host_decoded[0]:
            print(f"{n_host_decoded[0]} JPEG file(s) of a variant without a device path were decoded with cv2 on the host")
        if opt.save_txt:
            print(f"{totals[1]} labels saved to {label_dir} ({totals[2]} detections in {totals[0]} images)")
        print(f"Results saved to {save_dir}")
    return 0


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    return run(parse_opt(argv), argv)


if __name__ == "__main__":
    sys.exit(main())
