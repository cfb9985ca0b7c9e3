"""-m gpu: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded inputs."""
import numpy as np
import pytest

import flivo_parity as par
import flivo_scenes as scenes

pytestmark = pytest.mark.gpu


def test_knn_matches_bruteforce(pkg, small_world):
    pts = small_world["points"]
    rng = np.random.default_rng(3)
    q = pts[rng.choice(len(pts), 3000, replace=False)] + rng.normal(0, 0.05, (3000, 3)).astype(np.float32)
    q = np.concatenate([q, rng.uniform(-60, 60, (500, 3)).astype(np.float32)])   # far / sparse queries too
    with pkg.Flivo(downsample_size=small_world["ds"], initial_points=len(pts)) as F:
        F.Build(pts)
        for max_dist in (np.inf, 5.0, 0.3):
            gid, gsq, gxyz, gcnt = F.Nearest_Search(q, 5, max_dist)
            oid, osq, oxyz, ocnt, ties = par.oracle_knn(pts, q, 5, max_dist)
            assert ties == 0
            assert np.array_equal(gcnt, ocnt)
            assert np.array_equal(gid, oid)
            assert np.array_equal(gsq, osq)
            assert np.array_equal(gxyz, oxyz)


@pytest.mark.parametrize("mode", [0, 1])
def test_update_matches_oracle(pkg, small_world, small_scan, mode):
    par.check_update_against_oracle(pkg, small_world, small_scan, mode=mode, max_iter=4, verbose=True)


def test_update_sorted_scan_and_force_search(pkg, small_world, small_scan):
    par.check_update_against_oracle(pkg, small_world, small_scan, mode=0, max_iter=4, force_search=True,
                                    sort_scan=True, verbose=True)
