"""Golden vectors for the clustering test: runs the reference's OWN get_clusters source
(/root/reference/src/cluster_facilities.py, extracted with ast so its geopandas / shapely imports are not needed) with the
genuine sklearn.cluster.DBSCAN on seeded point sets.  Build container only (needs /root/reference):
python tools/make_golden_dbscan.py -> tests/golden/dbscan_ref.npz (scikit-learn 1.9.0 here; the reference pins 1.3.0)."""
import ast
import os

import numpy as np
from sklearn.cluster import DBSCAN

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
mod = ast.parse(open("/root/reference/src/cluster_facilities.py").read())
wanted = [n for n in mod.body if isinstance(n, ast.FunctionDef) and n.name == "get_clusters"]
ns = {"DBSCAN": DBSCAN, "np": np}
exec(compile(ast.Module(body=wanted, type_ignores=[]), "reference:src/cluster_facilities.py", "exec"), ns)
get_clusters = ns["get_clusters"]

rng = np.random.default_rng(3)
cases = {}


def farms(n_farms, cages, spread, noise, pitch=25.0):
    """Cage centres: rows of cages `pitch` m apart inside farms scattered over a few km, plus stray detections."""
    pts = []
    for _ in range(n_farms):
        c = rng.uniform(0, 4000, 2)
        k = int(rng.integers(2, cages))
        grid = np.stack(np.meshgrid(np.arange(k), np.arange(2)), -1).reshape(-1, 2) * pitch
        pts.append(c + grid + rng.normal(0, spread, grid.shape))
    pts.append(rng.uniform(0, 4000, (noise, 2)))
    return np.concatenate(pts)


cases["farms_default"] = (farms(12, 9, 1.0, 30, pitch=8.0), 10.0, 5)          # the reference's defaults: 10 m, 5 cages
cases["farms_eps30"] = (farms(20, 12, 3.0, 60), 30.0, 5)
cases["farms_eps50_min3"] = (farms(15, 8, 6.0, 40), 50.0, 3)
cases["min_samples_1"] = (rng.uniform(0, 100, (200, 2)), 5.0, 1)
cases["all_noise"] = (rng.uniform(0, 1e5, (50, 2)), 1.0, 3)
cases["one_point"] = (np.array([[3.0, 4.0]]), 10.0, 1)
cases["duplicates"] = (np.repeat(rng.uniform(0, 50, (20, 2)), 4, 0), 0.5, 4)
# integer lattice with eps equal to the spacing: distances exactly eps (inclusive) and contested border points
lat = np.stack(np.meshgrid(np.arange(12.0), np.arange(7.0)), -1).reshape(-1, 2)
lat = lat[rng.random(len(lat)) < 0.7]
cases["lattice_exact_eps"] = (lat * 10.0, 10.0, 4)
cases["lattice_diag"] = (lat * 3.0, 3.0 * 2 ** 0.5, 5)
# two dense blobs bridged by border points both can reach
a = rng.normal([0, 0], 0.5, (30, 2)); b = rng.normal([6, 0], 0.5, (30, 2))
cases["contested_border"] = (np.concatenate([a, b, [[3.0, 0.0]], [[3.0, 0.4]]]), 2.2, 6)
cases["shuffled_big"] = (rng.permutation(farms(60, 14, 1.5, 400, pitch=9.0)), 12.0, 5)
out = {}
for k, (pts, eps, ms) in cases.items():
    out[k + "_pts"] = pts.astype(np.float64)
    out[k + "_par"] = np.array([eps, ms], np.float64)
    out[k + "_lab"] = get_clusters(pts, eps, ms).astype(np.int64)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "dbscan_ref.npz"), **out)
print({k: (len(v[0]), int(out[k + "_lab"].max()) + 1, int((out[k + "_lab"] < 0).sum())) for k, v in cases.items()})
