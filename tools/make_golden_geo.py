"""Golden vectors for the geocode / area arithmetic: runs the reference's OWN source of convert_pix_to_m_bboxes
(/root/reference/src/process_yolo/geocode_results.py) and get_circle_area_from_bbox / get_square_area_from_bbox
(/root/reference/src/process_yolo/calc_net_areas.py), extracted with ast so their geopandas / pyproj imports are not
needed.  Build container only: python tools/make_golden_geo.py -> tests/golden/geo_ref.npz"""
import ast
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def extract(path, names, ns):
    mod = ast.parse(open(path).read())
    body = [n for n in mod.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert len(body) == len(names), (path, names)
    exec(compile(ast.Module(body=body, type_ignores=[]), "reference:" + path, "exec"), ns)


ns = {"os": os, "np": np, "LARGE_TIF_SIZE": 1024 * 6}
extract("/root/reference/src/process_yolo/geocode_results.py", ["convert_pix_to_m_bboxes"], ns)
extract("/root/reference/src/process_yolo/calc_net_areas.py", ["get_circle_area_from_bbox", "get_square_area_from_bbox"], ns)


class _Geom:
    def __init__(self, b):
        self.bounds = tuple(b)


class _Loc:
    def __init__(self, boxes):
        self.boxes = boxes

    def __getitem__(self, key):
        ind, col = key
        assert col == "geometry"
        return _Geom(self.boxes[ind])


class _Frame:  # stands in for the GeoDataFrame of download boxes: .loc[int, 'geometry'].bounds
    def __init__(self, boxes):
        self.loc = _Loc(boxes)


rng = np.random.default_rng(11)
boxes = {i: (x0, y0, x0 + 6144 * s, y0 + 6144 * s) for i, (x0, y0, s) in
         enumerate([(-500000.0, 6000000.0, 0.2), (123456.789, 5432109.876, 0.19999), (-1.5e5, 6.2e6, 0.2003)])}
frame = _Frame(boxes)
pts = []
for k in range(200):
    ind = int(rng.integers(0, 3)); xo = int(rng.integers(0, 6)) * 1024; yo = int(rng.integers(0, 6)) * 1024
    x = int(rng.integers(0, 1025)); y = int(rng.integers(0, 1025))
    xm, ym = ns["convert_pix_to_m_bboxes"](x, y, f"/a/b/ORTHOIMAGERY.ORTHOPHOTOS2020_{ind}_{xo}_{yo}.txt", frame)
    pts.append((ind, xo, yo, x, y, xm, ym))
areas = []
for k in range(100):
    w, h = float(rng.uniform(1, 60)), float(rng.uniform(1, 60))
    for xb in (False, True):
        for yb in (False, True):
            areas.append((0, w, h, xb, yb) + tuple(float(v) for v in ns["get_circle_area_from_bbox"](w, h, xb, yb)))
    areas.append((1, w, h, False, False) + tuple(float(v) for v in ns["get_square_area_from_bbox"](w, h)))
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "geo_ref.npz"), boxes=np.array([boxes[i] for i in range(3)]),
                    pts=np.array(pts, dtype=np.float64), areas=np.array(areas, dtype=np.float64))
print(len(pts), "points,", len(areas), "area rows")
