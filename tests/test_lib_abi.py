"""CPU: the C-ABI library loads, exports every symbol include/aqua_b200.h declares, refuses to run without a
GPU (no CPU fallback), and its host-side label formatter is bit-identical to the oracle's text."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from aquaculture_b200 import _lib, engine
from oracle import labels as OL

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "aqua_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(aq_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/aqua_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert lib.aq_abi_version() == 5


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    with pytest.raises(_lib.AquaError, match="no CUDA device|CPU path"):
        engine.Context(0)


def test_missing_library_is_loud(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libaqua_b200.so")
    with pytest.raises(_lib.AquaError, match="not built"):
        _lib.load()


@pytest.mark.parametrize("hw0", [(640, 640), (1024, 1024), (600, 1000), (333, 517)])
def test_label_text_matches_oracle(hw0):
    g = torch.Generator().manual_seed(hw0[0])
    n = 200
    net = (640, 640)
    xy = torch.rand(n, 2, generator=g) * 700 - 30
    wh = torch.rand(n, 2, generator=g) * 150
    det = torch.cat((xy - wh / 2, xy + wh / 2, torch.rand(n, 1, generator=g), torch.randint(0, 5, (n, 1), generator=g).float()), 1)
    det = det[det[:, 4].argsort(descending=True)]
    for save_conf in (True, False):
        exp = "".join(l + "\n" for l in OL.label_lines(det, net, hw0, save_conf))
        got = engine.format_label_text(det.numpy(), net, hw0, save_conf)
        assert got == exp


def test_label_text_half_pixel_rounding():
    # boxes exactly on .5 after rescale exercise round-half-even
    det = torch.tensor([[0.3125, 0.9375, 1.5625, 2.1875, 0.5, 1.0], [10.0, 10.0, 20.0, 20.0, 0.25, 0.0]])
    exp = "".join(l + "\n" for l in OL.label_lines(det, (640, 640), (1024, 1024)))
    assert engine.format_label_text(det.numpy(), (640, 640), (1024, 1024)) == exp


def test_write_labels_files(tmp_path):
    dets = np.zeros((3, 5, 6), np.float32)
    dets[0, 0] = [10, 10, 50, 50, 0.9, 1]
    dets[0, 1] = [100, 100, 150, 160, 0.6, 0]
    dets[2, 0] = [5, 5, 9, 9, 0.3, 2]
    counts = np.array([2, 0, 1], np.int32)
    paths = [str(tmp_path / f"L2020_{i}_0_0.txt") for i in range(3)]
    n = engine.write_labels(engine.Detections(dets, counts), (640, 640), [(640, 640)] * 3, paths)
    assert n == 2 and not os.path.exists(paths[1])
    a = np.loadtxt(paths[0])
    assert a.shape == (2, 6) and a[0, 5] < a[1, 5]          # ascending confidence
    assert np.loadtxt(paths[2]).shape == (6,)
    engine.write_labels(engine.Detections(dets, counts), (640, 640), [(640, 640)] * 3, paths)
    assert np.loadtxt(paths[0]).shape == (4, 6)              # append mode, as upstream


# ---------------------------------------------------------------- geocode + area arithmetic (host C)
def _geo_golden():
    import os
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "geo_ref.npz"))


def test_geocode_oracle_matches_reference_source():
    """oracle/geocode.py::pix_to_m vs the reference's own convert_pix_to_m_bboxes source (tests/golden/geo_ref.npz)."""
    from oracle import geocode as OG
    g = _geo_golden()
    for ind, xo, yo, x, y, xm, ym in g["pts"]:
        got = OG.pix_to_m(int(x), int(y), int(xo), int(yo), tuple(g["boxes"][int(ind)]))
        assert got == (xm, ym)
    for kind, w, h, xb, yb, est, var, lo, hi in g["areas"]:
        got = OG.square_area(w, h) if kind else OG.circle_area(w, h, bool(xb), bool(yb))
        assert tuple(float(v) for v in got) == (est, var, lo, hi)


def test_cage_areas_c_vs_reference_golden():
    g = _geo_golden()
    for kind, w, h, xb, yb, est, var, lo, hi in g["areas"]:
        assert engine.cage_areas(bool(kind), w, h, bool(xb), bool(yb)) == (est, var, lo, hi)


def test_geocode_label_text_c_vs_oracle():
    """aq_geocode_label_text on label text the C formatter wrote == np.loadtxt + the oracle's restatement of the
    reference loop, bit for bit (ints and float64), for one- and many-line files."""
    import io
    from oracle import geocode as OG
    g = _geo_golden()
    rng = np.random.default_rng(3)
    for trial in range(6):
        n = 1 if trial == 0 else int(rng.integers(2, 40))
        xy = rng.uniform(0, 600, (n, 2)).astype(np.float32)
        wh = rng.uniform(2, 80, (n, 2)).astype(np.float32)
        conf = np.sort(rng.uniform(0.25, 1, n).astype(np.float32))[::-1]
        det = np.concatenate([xy, xy + wh, conf[:, None], rng.integers(0, 3, (n, 1)).astype(np.float32)], 1).astype(np.float32)
        text = engine.format_label_text(np.ascontiguousarray(det), (640, 640), (1024, 1024))
        rows = np.loadtxt(io.StringIO(text))
        rows = [rows] if rows.ndim == 1 else list(rows)
        ind = trial % 3
        xo, yo = 1024 * (trial % 6), 1024 * ((trial * 5) % 6)
        exp = OG.geocode_rows(rows, xo, yo, tuple(g["boxes"][ind]))
        got = engine.geocode_label_text(text, xo, yo, g["boxes"][ind])
        assert got == exp
    with pytest.raises(engine.AquaError):  # five columns (no --save-conf): the consumer cannot use the file
        engine.geocode_label_text("0 0.5 0.5 0.1 0.1\n", 0, 0, g["boxes"][0])


def test_geocode_dets_equals_the_label_file_route():
    """aq_geocode_dets on a batch's detections == writing the label files, np.loadtxt, and the reference loop restated by
    the oracle (pinned by the reference's own source) -- ints and float64 bit for bit -- plus year / class name / order;
    the text hop (one np.loadtxt + one cv.imread per file upstream) is gone, the values are not."""
    import io
    from oracle import geocode as OG
    g = _geo_golden()
    rng = np.random.default_rng(5)
    n_img, max_det = 7, 40
    dets = np.zeros((n_img, max_det, 6), np.float32)
    counts = np.zeros(n_img, np.int32)
    stems, meta = [], []
    for i in range(n_img):
        n = 0 if i == 3 else int(rng.integers(1, max_det))
        xy = rng.uniform(0, 600, (n, 2)).astype(np.float32)
        wh = rng.uniform(2, 80, (n, 2)).astype(np.float32)
        conf = np.sort(rng.uniform(0.25, 1, n).astype(np.float32))[::-1]
        dets[i, :n] = np.concatenate([xy, xy + wh, conf[:, None], rng.integers(0, 5, (n, 1)).astype(np.float32)], 1)
        counts[i] = n
        ind, xo, yo, year = i % 3, 1024 * (i % 6), 1024 * ((i * 5) % 6), 2000 + i
        stems.append(f"ORTHOIMAGERY.ORTHOPHOTOS{year}_{ind}_{xo}_{yo}")
        meta.append((ind, xo, yo, year))
    res = engine.Detections(dets, counts)
    got = engine.geocode_detections(res, (640, 640), [(1024, 1024)] * n_img, stems, g["boxes"])
    exp = []
    names = {0: "circle_farm", 1: "square_farm", 2: "triangle_farm", 3: "other_farm", 4: "rectangle_farm"}
    for i in range(n_img):
        if counts[i] == 0:
            continue
        text = engine.format_label_text(np.ascontiguousarray(dets[i, :counts[i]]), (640, 640), (1024, 1024))
        rows = np.loadtxt(io.StringIO(text))
        rows = [rows] if rows.ndim == 1 else list(rows)
        ind, xo, yo, year = meta[i]
        for r in OG.geocode_rows(rows, xo, yo, tuple(g["boxes"][ind])):
            a = engine.transform_3857_to_3035([[r["xmin_m"], r["ymax_m"]], [r["xmax_m"], r["ymin_m"]]])
            exp.append(dict(image=stems[i] + ".jpeg", xmin=r["xmin"], xmax=r["xmax"], ymin=r["ymin"], ymax=r["ymax"],
                            xmin_m=a[0, 0], xmax_m=a[1, 0], ymin_m=a[1, 1], ymax_m=a[0, 1], type=names[r["cls"]], year=year,
                            det_conf=r["conf"], geometry_3857=(r["xmin_m"], r["ymin_m"], r["xmax_m"], r["ymax_m"])))
    assert got == exp and len(got) == int(counts.sum())
    for bad in ("a_b", "L2020_0_0", "L2020_0_0_0_0", "L2020_9_0_0", "Lxx2y_0_0_0", "L2020_0_x_0"):  # upstream: ValueError / KeyError
        with pytest.raises(engine.AquaError):
            engine.geocode_detections(res, (640, 640), [(1024, 1024)] * n_img, [bad] * n_img, g["boxes"])


def test_3857_to_3035_published_epsg_test_point():
    """EPSG Guidance Note 7-2, Lambert Azimuthal Equal Area (method 9820), ETRS89-extended / LAEA Europe: 50 N 5 E ->
    E 3 962 799.45 m, N 2 999 718.85 m; the natural origin 52 N 10 E -> the false easting / northing.  The Pseudo-Mercator
    leg (method 1024) is inverted from its defining formulas.  (PROJ itself is absent here: this is the published pin.)"""
    import ctypes as C
    from aquaculture_b200 import _lib
    ll = np.array([[5.0, 50.0], [10.0, 52.0]])
    out = np.empty_like(ll)
    assert _lib.load().aq_lonlat_to_3035(ll.ctypes.data, 2, out.ctypes.data) == 0
    assert abs(out[0, 0] - 3962799.45) < 0.005 and abs(out[0, 1] - 2999718.85) < 0.005
    assert abs(out[1, 0] - 4321000.0) < 1e-6 and abs(out[1, 1] - 3210000.0) < 1e-6
    R = 6378137.0
    merc = np.stack([np.radians(ll[:, 0]) * R, R * np.log(np.tan(np.pi / 4 + np.radians(ll[:, 1]) / 2))], 1)
    got = engine.transform_3857_to_3035(merc)
    assert np.abs(got - out).max() < 1e-6
    # area is what the reference uses 3035 for: a 1 km x 1 km geographic cell near Brittany keeps its area to 1e-4 relative
    # under the equal-area projection (ellipsoidal cell area vs projected quadrilateral)
    lon0, lat0, d = -3.0, 48.0, 0.01
    cell = np.array([[lon0, lat0], [lon0 + d, lat0], [lon0 + d, lat0 + d], [lon0, lat0 + d]])
    q = np.empty_like(cell)
    _lib.load().aq_lonlat_to_3035(cell.ctypes.data, 4, q.ctypes.data)
    shoelace = 0.5 * abs(sum(q[i, 0] * q[(i + 1) % 4, 1] - q[(i + 1) % 4, 0] * q[i, 1] for i in range(4)))
    a, f = 6378137.0, 1 / 298.257222101
    e2 = 2 * f - f * f
    e = np.sqrt(e2)
    qf = lambda p: (1 - e2) * (np.sin(p) / (1 - e2 * np.sin(p) ** 2) - np.log((1 - e * np.sin(p)) / (1 + e * np.sin(p))) / (2 * e))
    true = a * a * np.radians(d) * (qf(np.radians(lat0 + d)) - qf(np.radians(lat0))) / 2
    assert abs(shoelace / true - 1) < 1e-4
