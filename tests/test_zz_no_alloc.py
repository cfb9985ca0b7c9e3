"""-m gpu, collected last on purpose: flivo_knn_noalloc (the batched Nearest_Search whose scratch is a static device array of
the library: no cudaMalloc in the call, no counterpart in the reference) against
flivo_knn through the C ABI — bit-exact, as every neighbour test here — on two calls in a row that share the scratch."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _check(F, q, k, max_dist):
    want = F.Nearest_Search(q, k, max_dist)
    got = F.Nearest_Search_noalloc(q, k, max_dist)
    for w, g, name in zip(want, got, ("ids", "sqd", "xyz", "count")):
        assert np.array_equal(w, g), name
    return want


def test_knn_noalloc_matches_knn_twice_in_a_row(pkg, small_world):
    pts = small_world["points"]
    rng = np.random.default_rng(21)
    q1 = pts[rng.choice(len(pts), 6000, replace=False)] + rng.normal(0, 0.05, (6000, 3)).astype(np.float32)
    q1 = np.concatenate([q1, rng.uniform(-60, 60, (300, 3)).astype(np.float32)])            # sparse / far queries too
    q2 = pts[rng.choice(len(pts), 1500, replace=False)] + rng.normal(0, 0.2, (1500, 3)).astype(np.float32)
    with pkg.Flivo(downsample_size=small_world["ds"], initial_points=len(pts)) as F:
        F.Build(pts)
        ids1 = _check(F, q1, 5, np.inf)[0]
        assert (ids1 >= 0).all()
        # second call: fewer queries, another k and bound, 16-byte records -> must not see anything of the first call
        q2_4 = np.zeros((len(q2), 4), np.float32); q2_4[:, :3] = q2
        _check(F, q2_4, 3, 0.3)
        _check(F, q1, 5, 5.0)
        # beyond the stated maximum: a clear error, nothing computed
        big = np.zeros((32768 + 1, 3), np.float32)
        with pytest.raises(pkg.FlivoError) as e:
            F.Nearest_Search_noalloc(big, 5, np.inf)
        assert e.value.code == pkg.E_CAPACITY and "FLIVO_KNN_NOALLOC_MAX_QUERIES" in str(e.value)
        _check(F, np.concatenate([q1, q1, q1, q1])[:32768 - 7], 5, np.inf)                     # near the limit


def test_knn_noalloc_ivox_mode(pkg, small_world):
    pts = small_world["points"][:20000]
    rng = np.random.default_rng(22)
    q = pts[rng.choice(len(pts), 2000, replace=False)] + rng.normal(0, 0.05, (2000, 3)).astype(np.float32)
    with pkg.Flivo(downsample_size=0.5, initial_points=len(pts), ivox_nearby=pkg.IVOX_NEARBY18) as F:
        F.AddPoints(pts)
        _check(F, q, 5, 25.0)
        _check(F, q[:100], 5, 25.0)
