"""CPU (-m "not gpu"): the restated faster_lio::IVox (oracle/ivox_oracle.cpp) against an independent
numpy restatement, its edge cases, map_incremental's add rule on hand-made cases, and — the pin — against the
reference's own include/ivox3d headers compiled unmodified behind stand-ins for Eigen / pcl / glog
(oracle/_ref/libivox_ref.so, oracle/ivox_ref_wrap.cpp).  The reference has no tests for this branch."""
import numpy as np
import pytest

from oracle import pyoracle as orc


def _round_half_away(x):
    return (np.sign(x) * np.floor(np.abs(x) + np.float32(0.5))).astype(np.int64)


def _numpy_ivox_knn(pts, q, res, nearby, k=5, max_range=5.0):
    """Brute force: candidates = points whose Pos2Grid key differs from the query's by an offset with at most
    {1:0, 7:1, 19:2, 27:3}[nearby] non-zero components; k nearest by (float32 distance, index)."""
    inv = np.float32(1.0 / res)
    pk = _round_half_away(pts * inv)
    qk = _round_half_away(q * inv)
    max_nz = {1: 0, 7: 1, 19: 2, 27: 3}[nearby]
    ids = np.full((len(q), k), -1, np.int32)
    sqd = np.full((len(q), k), np.inf, np.float32)
    cnt = np.zeros(len(q), np.int32)
    for i in range(len(q)):
        dk = pk - qk[i]
        ok = (np.abs(dk).max(1) <= 1) & ((dk != 0).sum(1) <= max_nz)
        idx = np.nonzero(ok)[0]
        if len(idx) == 0:
            cnt[i] = -1
            continue
        d = pts[idx] - q[i]
        d2 = (d[:, 0] * d[:, 0] + (d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2])).astype(np.float32)
        keep = d2.astype(np.float64) < max_range * max_range
        idx, d2 = idx[keep], d2[keep]
        if len(idx) == 0:
            cnt[i] = -1
            continue
        o = np.lexsort((idx, d2))[:k]
        cnt[i] = len(o)
        ids[i, :len(o)] = idx[o]
        sqd[i, :len(o)] = d2[o]
    return ids, sqd, cnt


@pytest.mark.parametrize("nearby", [1, 7, 19, 27])
def test_ivox_knn_matches_numpy(nearby):
    rng = np.random.default_rng(100 + nearby)
    pts = rng.uniform(-6, 6, (6000, 3)).astype(np.float32)
    pts[:, 2] = (0.05 * rng.standard_normal(6000)).astype(np.float32)     # a noisy ground plane: dense voxels
    q = pts[rng.integers(0, len(pts), 400)] + rng.normal(0, 0.05, (400, 3)).astype(np.float32)
    q = np.vstack([q, rng.uniform(-8, 8, (50, 3)).astype(np.float32)])   # some queries off the surface
    v = orc.OracleIVox(0.5, nearby)
    v.add_points(pts[:3000])
    v.add_points(pts[3000:])                                              # ids continue across calls
    ids, sqd, xyz, cnt, ties = v.knn(q, 5, 5.0)
    wi, ws, wc = _numpy_ivox_knn(pts, q.astype(np.float32), 0.5, nearby)
    assert ties == 0
    assert np.array_equal(cnt, wc)
    assert np.array_equal(ids, wi)
    assert np.array_equal(sqd, ws)
    got = ids >= 0
    assert np.array_equal(xyz[got], pts[ids[got]])
    assert (cnt == -1).any() and (cnt == 5).any()


def test_ivox_max_range_and_small_k():
    pts = np.array([[0, 0, 0], [0.1, 0, 0], [0.2, 0, 0], [0.3, 0, 0]], np.float32)
    v = orc.OracleIVox(1.0, 27)
    v.add_points(pts)
    ids, sqd, _, cnt, _ = v.knn(np.array([[0.0, 0, 0]], np.float32), 3, 0.25)   # strict d < 0.0625: 0.0, 0.01, 0.04
    assert cnt[0] == 3 and list(ids[0]) == [0, 1, 2]
    ids, sqd, _, cnt, _ = v.knn(np.array([[0.0, 0, 0]], np.float32), 5, 0.2)    # 0.04 is not < 0.04 (float 0.2f^2 > 0.04?)
    assert cnt[0] in (2, 3)
    ids, _, _, cnt, _ = v.knn(np.array([[9.0, 9, 9]], np.float32), 5, 5.0)
    assert cnt[0] == -1 and (ids[0] == -1).all()


def test_ivox_pos2grid_rounds_half_away_from_zero():
    # keys: round(-0.5) = -1, round(0.5) = 1, round(1.5) = 2 (std::round), resolution 1
    pts = np.array([[-0.5, 0, 0], [0.5, 0, 0], [1.5, 0, 0]], np.float32)
    v = orc.OracleIVox(1.0, 1)     # CENTER: only the query's own voxel
    v.add_points(pts)
    assert v.num_grids() == 3
    ids, _, _, cnt, _ = v.knn(np.array([[-1.2, 0, 0], [1.2, 0, 0], [2.2, 0, 0], [0.2, 0, 0]], np.float32), 5, 5.0)
    assert list(ids[:, 0]) == [0, 1, 2, -1] and list(cnt) == [1, 1, 1, -1]


def test_ivox_lru_eviction_restated():
    # capacity 4: the 4th new voxel triggers the eviction of the least recently touched one (ivox3d.h:262-265)
    v = orc.OracleIVox(1.0, 1, capacity=4)
    v.add_points(np.array([[0, 0, 0], [5, 0, 0], [10, 0, 0]], np.float32))
    v.add_points(np.array([[0.1, 0, 0]], np.float32))           # touches voxel 0: now most recent
    v.add_points(np.array([[15, 0, 0]], np.float32))            # 4th voxel -> size >= capacity -> evict voxel of x=5
    assert v.evicted() == 1 and v.num_grids() == 3
    _, _, _, cnt, _ = v.knn(np.array([[5, 0, 0], [0, 0, 0], [10, 0, 0], [15, 0, 0]], np.float32), 5, 5.0)
    assert list(cnt) == [-1, 2, 1, 1]


def test_ivox_map_filter_rule():
    fs = 0.5
    # point in voxel [0,0.5)^3, centre (0.25,0.25,0.25)
    pw = np.array([[0.30, 0.30, 0.30]] * 5, np.float32)
    near = np.zeros((5, 5, 3), np.float32)
    cnt = np.zeros(5, np.int32)
    # 0: no cached neighbours -> add
    # 1: nearest neighbour farther than fs/2 from the centre on every axis -> no-need-downsample list
    near[1, :] = [0.9, 0.9, 0.9]; cnt[1] = 5
    # 2: five neighbours, one of them closer to the centre than the point -> dropped
    near[2, :] = [0.6, 0.3, 0.3]; near[2, 3] = [0.26, 0.25, 0.25]; cnt[2] = 5
    # 3: five neighbours, all farther from the centre than the point -> add
    near[3, :] = [0.6, 0.3, 0.3]; cnt[3] = 5
    # 4: fewer than five neighbours (and near[0] not far on every axis) -> add whatever their distance
    near[4, 0] = [0.25, 0.25, 0.25]; cnt[4] = 2
    flags = orc.ivox_map_filter(pw, near, cnt, fs)
    assert list(flags) == [1, 2, 0, 1, 1]


def test_gnss_block_matches_numpy():
    """Restated GNSSProcessing::computeH (src/GNSS_Processing.cpp:204-237) against the matrix formula in numpy."""
    rng = np.random.default_rng(7)
    a = rng.normal(0, 0.4, 3)
    th = np.linalg.norm(a)
    Kx = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(th) / th * Kx + (1 - np.cos(th)) / th ** 2 * Kx @ Kx
    pos = rng.normal(0, 5, 3)
    lever = np.array([0.1, -0.2, 0.5])
    std = np.array([0.02, 0.03, 0.05])
    gpos = pos - R @ lever + np.array([0.3, -0.1, 0.2])
    st = orc.State.make(R=R, t=pos)
    H, L = orc.gnss_block(st, gpos, lever, std)
    S = np.array([[0, -lever[2], lever[1]], [lever[2], 0, -lever[0]], [-lever[1], lever[0], 0]])
    Hsub = np.hstack([R @ S, -np.eye(3)])
    W = np.diag([1 / std[0], 1 / std[1], 100 / std[2]])
    err = gpos - pos + R @ lever
    assert np.allclose(H, Hsub.T @ W @ Hsub, rtol=1e-13, atol=1e-12)
    assert np.allclose(L, Hsub.T @ W @ err, rtol=1e-13, atol=1e-12)
    # observation farther than 2 m from the state: ignored (:221-222)
    H, L = orc.gnss_block(st, gpos + np.array([2.5, 0, 0]), lever, std)
    assert not H.any() and not L.any()


# ---- the restatement against the reference's own header (oracle/_ref/libivox_ref.so) -------------------------------
def _rows_as_sets(xyz, cnt):
    """Each row's points sorted lexicographically (the reference returns them in nth_element order)."""
    out = []
    for i in range(len(cnt)):
        c = max(int(cnt[i]), 0)
        r = xyz[i, :c]
        out.append(r[np.lexsort((r[:, 2], r[:, 1], r[:, 0]))] if c else r)
    return out


needs_ref = pytest.mark.skipif(not orc.have_ivox_ref(), reason="oracle/_ref/libivox_ref.so not built (reference tree absent)")


@needs_ref
@pytest.mark.parametrize("nearby", [1, 7, 19, 27])
def test_restatement_matches_the_compiled_reference_header(nearby):
    """Same AddPoints batches, same queries: voxel count, found/not-found, neighbour count and the neighbour SET of every
    query must be those of faster_lio::IVox itself (include/ivox3d/ivox3d.h compiled unmodified)."""
    rng = np.random.default_rng(100 + nearby)
    res = 0.5
    pts = (rng.random((60000, 3)) * np.array([24.0, 24.0, 6.0]) - np.array([12.0, 12.0, 3.0])).astype(np.float32)
    # points exactly half-way between voxel centres (p / res = n + 0.5) and on negative coordinates: the rounding rule
    edge = (np.round(rng.random((500, 3)) * 40 - 20) * res + res / 2).astype(np.float32)
    edge[:, 1:] += (rng.random((500, 2)) * 0.2).astype(np.float32)      # (x stays on the boundary; no two points coincide)
    edge = np.unique(edge, axis=0)
    O, R = orc.OracleIVox(res, nearby), orc.RefIVox(res, nearby)
    for batch in (pts[:25000], edge, pts[25000:]):
        O.add_points(batch)
        R.add_points(batch)
    assert O.num_grids() == R.num_grids()
    q = (rng.random((5000, 3)) * np.array([30.0, 30.0, 9.0]) - np.array([15.0, 15.0, 4.5])).astype(np.float32)
    q[:200] = edge[:200] + (rng.random((200, 3)) * 0.05).astype(np.float32)
    for k, max_range in ((5, 5.0), (3, 5.0), (1, 5.0), (5, 0.3)):
        oid, osq, oxyz, ocnt, ties = O.knn(q, k, max_range)
        rxyz, rcnt = R.knn(q, k, max_range)
        assert ties == 0
        assert np.array_equal(ocnt, rcnt), (k, max_range)
        assert (ocnt == -1).any() and (ocnt == k).any()
        a, b = _rows_as_sets(oxyz, ocnt), _rows_as_sets(rxyz, rcnt)
        bad = [i for i in range(len(q)) if not np.array_equal(a[i], b[i])]
        assert not bad, (k, max_range, len(bad), bad[:5])
        # element 0 of the reference's list is the nearest point (the final nth_element at position 0, ivox3d.h:178)
        has = ocnt > 0
        assert np.array_equal(rxyz[has, 0], oxyz[has, 0])


@needs_ref
def test_lru_eviction_matches_the_compiled_reference_header():
    """capacity_ voxels (ivox3d.h:262-265): a new voxel that fills the map evicts the least recently TOUCHED one."""
    rng = np.random.default_rng(7)
    res, cap = 0.5, 300
    O, R = orc.OracleIVox(res, 19, cap), orc.RefIVox(res, 19, cap)
    for b in range(12):
        centre = rng.random(3) * 10 - 5
        batch = (centre + rng.normal(0, 1.5, (800, 3))).astype(np.float32)
        O.add_points(batch)
        R.add_points(batch)
        assert O.num_grids() == R.num_grids() and O.num_grids() < cap
    assert O.evicted() > 0
    q = (rng.random((3000, 3)) * 14 - 7).astype(np.float32)
    oid, osq, oxyz, ocnt, ties = O.knn(q, 5, 5.0)
    rxyz, rcnt = R.knn(q, 5, 5.0)
    assert ties == 0 and np.array_equal(ocnt, rcnt)
    a, b = _rows_as_sets(oxyz, ocnt), _rows_as_sets(rxyz, rcnt)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))


# ---- the restatement against the committed outputs of the compiled reference header (runs without /root/reference) --
import os  # noqa: E402

import flivo_parity as par  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ivox_reference.npz")


@pytest.mark.parametrize("nearby", [1, 7, 19, 27])
def test_restatement_matches_reference_golden(nearby):
    gold = np.load(GOLD)
    O = orc.OracleIVox(float(gold["resolution"]), nearby)

    def knn(q, k, max_range):
        _, _, xyz, cnt, ties = O.knn(q, k, max_range)
        assert ties == 0
        return xyz, cnt
    par.check_ivox_golden(gold, nearby, O.add_points, O.num_grids, knn)


def test_restated_lru_matches_reference_golden():
    gold = np.load(GOLD)
    O = orc.OracleIVox(float(gold["resolution"]), 19, int(gold["lru_capacity"]))
    for b in gold["lru_batches"]:
        O.add_points(b)
    assert O.num_grids() == int(gold["lru_grids"]) and O.evicted() > 0
    _, _, xyz, cnt, ties = O.knn(gold["lru_queries"], 5, 5.0)
    assert ties == 0 and np.array_equal(cnt, gold["lru_cnt"])
    assert np.array_equal(par.lexsorted_rows(xyz, cnt), gold["lru_set"])
