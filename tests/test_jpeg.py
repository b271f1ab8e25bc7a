"""JPEG tiles.  CPU: the host Huffman parser of the library + the oracle's integer restatement of the
reconstruction must reproduce cv2.imdecode (the genuine libjpeg-turbo behind upstream's cv2.imread) bit for bit on the
committed golden files and on fresh files; unsupported variants are refused.  GPU: aq_jpeg_reconstruct against the same
vectors, and detections from JPEG tiles identical to detections from the cv2-decoded pixels."""
import io
import os

import cv2
import numpy as np
import pytest
import torch

from aquaculture_b200 import _lib, engine, jpeg, weights
from oracle import jpeg as OJ

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "jpeg_ref.npz"))
CASES = sorted(k[4:] for k in GOLD.files if k.startswith("jpg_"))


def host_decode(blob):
    info = jpeg.probe(blob)
    coef = np.zeros(info.coef_count, np.int16)
    quant = np.zeros(192, np.uint16)
    jpeg.entropy_decode(blob, coef, quant)
    return info, coef, quant


def tile_scene(h, w, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, (h // 16 + 2, w // 16 + 2, 3), dtype=np.uint8)
    a = cv2.resize(a, (w, h), interpolation=cv2.INTER_CUBIC).astype(np.int32) + rng.integers(-12, 12, (h, w, 3))
    return np.clip(a, 0, 255).astype(np.uint8)


@pytest.mark.parametrize("case", CASES)
def test_oracle_and_host_parser_match_golden(case):
    blob = GOLD["jpg_" + case].tobytes()
    info, coef, quant = host_decode(blob)
    ref = GOLD["bgr_" + case]
    assert (info.height, info.width) == ref.shape[:2]
    got = OJ.reconstruct(coef, quant, info.width, info.height, info.ncomp, info.hs)
    assert np.array_equal(got, ref)


def test_golden_still_matches_this_cv2():
    """The committed vectors were produced by cv2.imdecode; the library in this environment must agree with them
    (guards the oracle against a cv2 build with different IDCT / upsampling defaults)."""
    for case in CASES[::5]:
        assert np.array_equal(cv2.imdecode(GOLD["jpg_" + case], cv2.IMREAD_COLOR), GOLD["bgr_" + case])


@pytest.mark.parametrize("hw,quality", [((640, 640), 75), ((1024, 1024), 75), ((200, 328), 90)])
def test_fresh_tiles_match_cv2(hw, quality):
    img = tile_scene(*hw, seed=hw[0] + quality)
    ok, enc = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, quality])
    info, coef, quant = host_decode(enc.tobytes())
    assert (info.ncomp, info.hs, info.vs) == (3, 2, 2)
    got = OJ.reconstruct(coef, quant, info.width, info.height, info.ncomp, info.hs)
    assert np.array_equal(got, cv2.imdecode(enc, cv2.IMREAD_COLOR))


@pytest.mark.parametrize("case", CASES)
def test_packed_form_equals_dense(case):
    blob = GOLD["jpg_" + case].tobytes()
    info, coef, quant = host_decode(blob)
    desc = np.zeros(info.coef_count // 64, np.uint32)
    ent = np.zeros(info.coef_count, np.uint32)
    q2 = np.zeros(192, np.uint16)
    n = jpeg.entropy_decode_packed(blob, desc, ent, q2)
    assert n == int((desc & 127).sum()) == int(np.count_nonzero(coef))
    assert np.array_equal(jpeg.unpack(desc, ent), coef) and np.array_equal(q2, quant)


@pytest.mark.parametrize("sub_bits", [32, 256, 1024])
@pytest.mark.parametrize("case", [c for c in CASES if "rst" not in c])
def test_parallel_huffman_schedule_equals_sequential(case, sub_bits):
    """The device's decoding schedule (guess, relax, scan, write, DC scan) rehearsed on the CPU with the same core."""
    blob = GOLD["jpg_" + case].tobytes()
    info, coef, quant = host_decode(blob)
    info2, scan, nbytes, tables, q2 = jpeg.scan_prepare(blob)
    got, rounds = jpeg.debug_parallel_decode(info2, scan, nbytes, tables, sub_bits)
    assert np.array_equal(got, coef) and np.array_equal(q2, quant) and rounds >= 1


def test_parallel_huffman_on_a_tile():
    ok, enc = cv2.imencode(".jpg", tile_scene(1024, 1024, seed=3), [cv2.IMWRITE_JPEG_QUALITY, 75])
    info, coef, quant = host_decode(enc.tobytes())
    info2, scan, nbytes, tables, q2 = jpeg.scan_prepare(enc.tobytes())
    got, rounds = jpeg.debug_parallel_decode(info2, scan, nbytes, tables, 1024)
    assert np.array_equal(got, coef) and rounds < 64


def test_scan_prepare_batch_equals_per_tile(monkeypatch):
    """The one-call batch preparation (C threads, straight into staging) writes what the per-tile call writes."""
    blobs = [cv2.imencode(".jpg", tile_scene(200, 328, seed=i), [cv2.IMWRITE_JPEG_QUALITY, 60 + 7 * i])[1].tobytes() for i in range(5)]

    class FakeDet:  # staging without a GPU: plain numpy instead of pinned memory
        batch = 8

        def jpeg_scan_staging(self, n, min_bytes=0):
            return (np.zeros(min_bytes + 64, np.uint32).view(np.uint8)[:min_bytes + 64], np.zeros(n, np.uint64), np.zeros(n, np.uint32),
                    np.zeros((n, _lib.AQ_JPEG_TABLE_BYTES), np.uint8), np.zeros((n, 192), np.uint16), np.zeros(n, np.int32))

    (sc, off, ln, tab, q, stt), total, i0 = engine.prepare_jpeg_batch(FakeDet(), blobs, n_threads=3)
    assert (i0.width, i0.height) == (328, 200) and total <= sc.size
    for i, b in enumerate(blobs):
        info, scan, nbytes, tables, quant = jpeg.scan_prepare(b)
        assert ln[i] == nbytes and off[i] % 4 == 0
        assert np.array_equal(sc[int(off[i]):int(off[i]) + scan.size], scan)
        assert np.array_equal(tab[i], tables) and np.array_equal(q[i], quant)
    with pytest.raises(jpeg.JpegUnsupported):
        engine.prepare_jpeg_batch(FakeDet(), blobs[:2] + [GOLD["unsupported_progressive"].tobytes()])


def test_restart_interval_files_are_left_to_the_host_decoder():
    case = next(c for c in CASES if "rst" in c and c.startswith("96x128"))
    with pytest.raises(jpeg.JpegUnsupported):
        jpeg.scan_prepare(GOLD["jpg_" + case].tobytes())


@pytest.mark.parametrize("name", ["unsupported_progressive", "unsupported_422", "unsupported_cmyk"])
def test_unsupported_variants_are_refused(name):
    with pytest.raises(jpeg.JpegUnsupported):
        jpeg.probe(GOLD[name].tobytes())
    with pytest.raises(jpeg.JpegUnsupported):
        jpeg.entropy_decode(GOLD[name].tobytes(), np.zeros(1 << 16, np.int16), np.zeros(192, np.uint16))


def test_corrupt_and_truncated_files_error():
    blob = GOLD["jpg_" + CASES[0]].tobytes()
    with pytest.raises(_lib.AquaError):
        jpeg.probe(b"not a jpeg at all")
    with pytest.raises(_lib.AquaError):
        jpeg.probe(blob[:40])                       # cut inside the headers
    info = jpeg.probe(blob)
    with pytest.raises((_lib.AquaError, AssertionError)):
        jpeg.entropy_decode(blob, np.zeros(info.coef_count - 64, np.int16), np.zeros(192, np.uint16))  # buffer too small
    # a scan cut short decodes like libjpeg does (zero bits fed after the end of data): no crash, same geometry
    coef = np.zeros(info.coef_count, np.int16)
    try:
        jpeg.entropy_decode(blob[:len(blob) // 2], coef, np.zeros(192, np.uint16))
    except _lib.AquaError:
        pass


def test_oversubscribed_dht_is_rejected_not_overrun():
    """A DHT whose code-length counts over-subscribe the code space (bits[1] = 255) used to index far past the 512-entry
    lookup table while it was being filled (stack overflow): it must be refused before anything is written."""
    for bits in ([255] + [0] * 15, [0, 5] + [0] * 14, [1, 1, 3] + [0] * 13, [0] * 8 + [255] + [0] * 7):
        n = min(sum(bits), 256)
        dht = bytes([0x00] + bits) + bytes(range(n))
        blob = b"\xff\xd8" + b"\xff\xc4" + (len(dht) + 2).to_bytes(2, "big") + dht + b"\xff\xd9"
        with pytest.raises(_lib.AquaError):
            jpeg.probe(blob)


def test_header_fuzz_never_crashes():
    """Random byte flips / truncations in the marker segments of valid files: every entry point either succeeds or returns
    an error status -- the parser runs on io threads inside the CLI, so a crash would take the whole run down."""
    rng = np.random.default_rng(7)
    for case in CASES[::6]:
        blob = np.array(GOLD["jpg_" + case])
        hdr = min(len(blob), 700)
        for _ in range(150):
            b = blob.copy()
            for _ in range(int(rng.integers(1, 6))):
                b[int(rng.integers(2, hdr))] = int(rng.integers(0, 256))
            if rng.random() < 0.3:
                b = b[:int(rng.integers(4, len(b)))].copy()
            try:
                info = jpeg.probe(b)
                if 0 < info.coef_count <= (1 << 24):
                    jpeg.entropy_decode(b, np.zeros(info.coef_count, np.int16), np.zeros(192, np.uint16))
                    jpeg.scan_prepare(b)
            except _lib.AquaError:
                pass


# ------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def ctx():
    c = engine.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("packed", [True, False])
def test_device_reconstruct_matches_golden(ctx, case, packed):
    got = engine.decode_jpeg_files(ctx, [GOLD["jpg_" + case].tobytes()], packed=packed)
    assert np.array_equal(got[0], GOLD["bgr_" + case])


@pytest.mark.gpu
@pytest.mark.parametrize("case", [c for c in CASES if "rst" not in c])
def test_device_huffman_matches_golden(ctx, case):
    got = engine.decode_jpeg_files(ctx, [GOLD["jpg_" + case].tobytes()], mode="device")
    assert np.array_equal(got[0], GOLD["bgr_" + case])


@pytest.mark.gpu
@pytest.mark.parametrize("hw,n,q0", [((1024, 1024), 6, 60), ((640, 640), 8, 70), ((200, 328), 3, 90), ((512, 512), 2, 100)])
def test_device_huffman_batch_matches_cv2(ctx, hw, n, q0):
    rng = np.random.default_rng(hw[0])
    blobs = []
    for i in range(n):
        img = tile_scene(*hw, seed=200 + i) if q0 < 100 else rng.integers(0, 256, (*hw, 3), dtype=np.uint8)  # q100 noise: slowest to fall into step
        ok, enc = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, min(100, q0 + 5 * i)])
        blobs.append(enc.tobytes())
    got = engine.decode_jpeg_files(ctx, blobs, mode="device")
    for i, b in enumerate(blobs):
        assert np.array_equal(got[i], cv2.imdecode(np.frombuffer(b, np.uint8), cv2.IMREAD_COLOR)), i


@pytest.mark.gpu
def test_device_huffman_reports_corrupt_scan(ctx):
    ok, enc = cv2.imencode(".jpg", tile_scene(256, 256, seed=1), [cv2.IMWRITE_JPEG_QUALITY, 75])
    blob = bytearray(enc.tobytes())
    del blob[len(blob) // 2:len(blob) - 2]   # half of the scan is missing
    with pytest.raises(_lib.AquaError, match="status"):
        engine.decode_jpeg_files(ctx, [bytes(blob)], mode="device")


@pytest.mark.gpu
def test_device_reconstruct_rgb_order(ctx):
    case = CASES[0]
    got = engine.decode_jpeg_files(ctx, [GOLD["jpg_" + case].tobytes()], bgr=False)
    assert np.array_equal(got[0], GOLD["bgr_" + case][:, :, ::-1])


@pytest.mark.gpu
@pytest.mark.parametrize("hw,n", [((1024, 1024), 5), ((640, 640), 8), ((200, 328), 3)])
def test_device_reconstruct_batch_matches_cv2(ctx, hw, n):
    blobs = []
    for i in range(n):
        ok, enc = cv2.imencode(".jpg", tile_scene(*hw, seed=100 + i), [cv2.IMWRITE_JPEG_QUALITY, 60 + 5 * i])  # tables differ per file
        blobs.append(enc.tobytes())
    got = engine.decode_jpeg_files(ctx, blobs)
    for i, b in enumerate(blobs):
        assert np.array_equal(got[i], cv2.imdecode(np.frombuffer(b, np.uint8), cv2.IMREAD_COLOR)), i


@pytest.mark.gpu
def test_detect_from_jpeg_equals_detect_from_pixels():
    """Whole path: JPEG tiles -> (host Huffman, device IDCT/colour) -> letterbox -> conv stack -> NMS gives exactly the
    detections of the same detector fed the cv2.imread pixels."""
    os.environ["AQUA_AUTOTUNE"] = "0"
    try:
        sd = weights.synthetic_state(3, 0)
        det = engine.TileDetector(sd, 3, batch=4, imgsz=(640, 640), device=0)
    finally:
        os.environ.pop("AQUA_AUTOTUNE", None)
    blobs, pix = [], []
    for i in range(4):
        ok, enc = cv2.imencode(".jpg", tile_scene(1024, 1024, seed=i), [cv2.IMWRITE_JPEG_QUALITY, 75])
        blobs.append(enc.tobytes())
        pix.append(cv2.imdecode(enc, cv2.IMREAD_COLOR))
    ref = det.detect(torch.from_numpy(np.stack(pix)).pin_memory(), bgr=True)
    ref_d, ref_c = ref.dets.copy(), ref.counts.copy()
    info = jpeg.probe(blobs[0])
    tl = [engine.entropy_decode_file(b) for b in blobs]
    for rep in range(3):  # both staging areas, and a reused one
        desc, ent, base, quant = det.jpeg_staging(4, info.coef_count, sum(t.entries.size for t in tl))
        total = engine.pack_jpeg_tiles(tl, desc, ent, base, quant)
        got = det.detect_wait(det.detect_submit_jpeg(desc, ent, base, total, quant, info.width, info.height, info.ncomp, info.hs))
        assert np.array_equal(got.counts, ref_c)
        for i in range(4):
            assert np.array_equal(got.dets[i, :ref_c[i]], ref_d[i, :ref_c[i]])
    sl = [engine.prepare_jpeg_file(b) for b in blobs]
    for rep in range(3):  # all-device decoder
        sc, off, ln, tab, q, stt = det.jpeg_scan_staging(4, sum(t.scan.size for t in sl))
        total = engine.pack_jpeg_scans(sl, sc, off, ln, tab, q)
        got = det.detect_wait(det.detect_submit_jpeg_scans(sc, off, ln, total, tab, q, stt, info.width, info.height, info.ncomp, info.hs))
        assert not stt.any() and np.array_equal(got.counts, ref_c)
        for i in range(4):
            assert np.array_equal(got.dets[i, :ref_c[i]], ref_d[i, :ref_c[i]])
    det.close()


@pytest.mark.gpu
def test_cli_device_jpeg_equals_host_decode(tmp_path, monkeypatch, capsys):
    """detect.py over a directory of .jpeg tiles (what the reference's tiler writes [REF src/load_data/tile_tifs.py:74]):
    the default --jpeg-decode device run writes exactly the label files of the cv2.imread run; a progressive file and a
    PNG in the same directory go through cv2 and are reported."""
    from PIL import Image
    from aquaculture_b200 import detect_cli as cli
    from aquaculture_b200 import tiles as T
    monkeypatch.setenv("AQUA_AUTOTUNE", "0")
    src = tmp_path / "jpegs"
    src.mkdir()
    parts = T.synthetic_tiles(range(7), 1024, 1024, seed=9)
    for i in range(5):
        extra = [cv2.IMWRITE_JPEG_RST_INTERVAL, 16] if i == 4 else []   # one file with restart markers
        cv2.imwrite(str(src / f"ORTHO2020_{i}_0_0.jpeg"), parts[i], [cv2.IMWRITE_JPEG_QUALITY, 75] + extra)
    Image.fromarray(parts[5]).save(str(src / "ORTHO2020_5_0_0.jpeg"), "JPEG", progressive=True)
    cv2.imwrite(str(src / "ORTHO2020_6_0_0.png"), parts[6])
    common = ["--weights", "synthetic", "--source", str(src), "--nosave", "--save-txt", "--save-conf", "--batch-size", "4",
              "--project", str(tmp_path / "runs")]
    assert cli.main(common + ["--name", "dev"]) == 0
    assert "1 JPEG file(s) of a variant without a device path" in capsys.readouterr().out
    assert cli.main(common + ["--name", "host", "--jpeg-decode", "host"]) == 0
    assert cli.main(common + ["--name", "hyb", "--jpeg-decode", "hybrid"]) == 0
    lhy = {p.name: p.read_text() for p in (tmp_path / "runs" / "hyb" / "labels").glob("*.txt")}
    ld = {p.name: p.read_text() for p in (tmp_path / "runs" / "dev" / "labels").glob("*.txt")}
    lh = {p.name: p.read_text() for p in (tmp_path / "runs" / "host" / "labels").glob("*.txt")}
    assert len(lh) >= 5 and ld == lh and lhy == lh


@pytest.mark.gpu
def test_cli_jpeg_raster_decoded_on_device(tmp_path, monkeypatch):
    """--raster-tile on a .jpeg raster: the raster is decoded on the GPU and its windows are cut there (rows f-1 + f-3
    together); labels equal those of the cv2.imread run, ragged right strip included."""
    from aquaculture_b200 import detect_cli as cli
    from aquaculture_b200 import tiles as T
    monkeypatch.setenv("AQUA_AUTOTUNE", "0")
    Tl = 512
    parts = T.synthetic_tiles(range(6), Tl, Tl, seed=12)
    raster = np.concatenate([np.concatenate(list(parts[0:3]), axis=1), np.concatenate(list(parts[3:6]), axis=1)], axis=0)
    raster = cv2.GaussianBlur(np.ascontiguousarray(raster[:, :Tl * 2 + 200]), (0, 0), 1.5)   # 1024 x 1224, compressible
    rdir = tmp_path / "raster"
    rdir.mkdir()
    cv2.imwrite(str(rdir / "BIG2020_7.jpeg"), raster, [cv2.IMWRITE_JPEG_QUALITY, 90])
    common = ["--weights", "synthetic", "--source", str(rdir), "--nosave", "--save-txt", "--save-conf", "--batch-size", "4",
              "--project", str(tmp_path / "runs"), "--raster-tile", str(Tl)]
    assert cli.main(common + ["--name", "dev"]) == 0
    assert cli.main(common + ["--name", "host", "--jpeg-decode", "host"]) == 0
    ld = {p.name: p.read_text() for p in (tmp_path / "runs" / "dev" / "labels").glob("*.txt")}
    lh = {p.name: p.read_text() for p in (tmp_path / "runs" / "host" / "labels").glob("*.txt")}
    assert len(lh) >= 1 and ld == lh
    got = engine.decode_jpeg_raster(engine.Context(0), np.fromfile(str(rdir / "BIG2020_7.jpeg"), np.uint8))
    assert np.array_equal(got.cpu().numpy(), cv2.imread(str(rdir / "BIG2020_7.jpeg")))
