"""facility clustering.  aq_dbscan (host C) must give exactly the labels of the reference's own get_clusters
[REF src/cluster_facilities.py:11-15] run with the genuine sklearn.cluster.DBSCAN (tests/golden/dbscan_ref.npz, made by
tools/make_golden_dbscan.py), and of the sklearn in this environment on fresh points."""
import os

import numpy as np
import pytest

from aquaculture_b200 import engine

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "dbscan_ref.npz"))
CASES = sorted(f[:-4] for f in GOLD.files if f.endswith("_pts"))


@pytest.mark.parametrize("case", CASES)
def test_labels_match_reference_get_clusters(case):
    eps, ms = GOLD[case + "_par"]
    got = engine.dbscan_labels(GOLD[case + "_pts"], eps, int(ms))
    assert np.array_equal(got, GOLD[case + "_lab"])


@pytest.mark.parametrize("seed", range(6))
def test_labels_match_sklearn_on_fresh_points(seed):
    from sklearn.cluster import DBSCAN
    rng = np.random.default_rng(seed)
    n = int(rng.integers(1, 1500))
    centres = rng.uniform(0, 500, (max(1, n // 40), 2))
    pts = centres[rng.integers(0, len(centres), n)] + rng.normal(0, 4.0, (n, 2))
    pts = np.round(pts, int(rng.integers(0, 3)))          # coarse coordinates: many exact ties at the threshold
    eps, ms = float(rng.choice([2.0, 5.0, 10.0])), int(rng.integers(1, 8))
    assert np.array_equal(engine.dbscan_labels(pts, eps, ms), DBSCAN(eps=eps, min_samples=ms).fit(pts).labels_)


def test_bad_arguments():
    with pytest.raises(engine.AquaError):
        engine.dbscan_labels(np.zeros((3, 2)), 0.0, 5)
    with pytest.raises(engine.AquaError):
        engine.dbscan_labels(np.array([[0.0, np.nan]]), 1.0, 1)
    assert engine.dbscan_labels(np.zeros((0, 2)), 1.0, 1).size == 0
