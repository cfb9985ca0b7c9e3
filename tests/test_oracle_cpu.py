"""CPU-only: the oracle (oracle/) against (a) golden vectors produced by the reference's own ikd_Tree.cpp
(tests/golden/make_golden.py), (b) the reference ikd-Tree itself when oracle/_ref is built, (c) known-answer and
numpy float64 cross-checks of the Eigen-dependent half (esti_plane, solves), whose parity is otherwise unpinned."""
import os

import numpy as np
import pytest

from oracle import pyoracle as orc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _sets_equal(a_ids, b_ids):
    return np.array_equal(np.sort(a_ids, axis=1), np.sort(b_ids, axis=1))


def test_knn_oracle_matches_ikdtree_golden():
    g = np.load(os.path.join(GOLD, "knn_ikdtree.npz"))
    for tag, md in (("inf", np.inf), ("5", 5.0), ("03", 0.3)):
        oid, osq, oxyz, ocnt, ties = orc.knn_bruteforce(g["map"], g["queries"], 5, md)
        assert ties == 0
        assert np.array_equal(ocnt, g["cnt_" + tag])
        assert np.array_equal(osq, g["sqd_" + tag])            # ascending float32 squared distances, bit-exact
        assert _sets_equal(oid, g["ids_" + tag])               # membership
        assert np.array_equal(oid, g["ids_" + tag])            # and order (no equal distances in this data)


def test_map_ops_oracle_matches_ikdtree_golden():
    g = np.load(os.path.join(GOLD, "map_ops_ikdtree.npz"))
    M = orc.OracleMap(float(g["ds"]))
    M.build(g["base"])
    for b in range(3):
        assert M.add_points(g["batch%d" % b], True) == int(g["ret%d" % b])
        xyz, ids = M.live()
        o = np.argsort(ids)
        assert np.array_equal(ids[o], g["live%d_ids" % b])
        assert np.array_equal(xyz[o], g["live%d_xyz" % b])
    assert M.add_points(g["batch_nods"], False) == int(g["ret_nods"])
    assert M.delete_boxes(g["boxes"]) == int(g["ret_del"])
    xyz, ids = M.live()
    o = np.argsort(ids)
    assert np.array_equal(ids[o], g["final_ids"]) and np.array_equal(xyz[o], g["final_xyz"])
    oid, osq, _, ocnt, ties = orc.knn_bruteforce(xyz, g["queries"], 5, np.inf, ids=ids)
    assert ties == 0
    assert np.array_equal(ocnt, g["q_cnt"]) and np.array_equal(osq, g["q_sqd"]) and np.array_equal(oid, g["q_ids"])


@pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built (needs /root/reference at build time)")
def test_oracle_matches_live_reference_ikdtree(small_world, small_scan):
    """Same checks against the reference tree run live, on the world / scan the GPU parity tests use."""
    pts = small_world["points"]
    T = orc.RefTree(small_world["ds"])
    T.build(pts)
    st = orc.State.from_array(small_scan["x_prior"])
    q = orc.body_to_world(small_scan["body"], st)
    rid, rsq, rxyz, rcnt = T.knn(q, 5, np.inf)
    oid, osq, oxyz, ocnt, ties = orc.knn_bruteforce(pts, q, 5, np.inf)
    assert ties == 0
    assert np.array_equal(rcnt, ocnt) and np.array_equal(rsq, osq) and np.array_equal(rid, oid) and np.array_equal(rxyz, oxyz)
    # whole iterated update: reference tree + OpenMP loop vs brute-force oracle loop
    import flivo_parity as par
    for mode in (orc.MODE_V3, orc.MODE_V1):
        npd = 18 if mode == orc.MODE_V1 else 23
        r = T.iekf_update(small_scan["body"], st, np.eye(npd) * 1e-4, mode=mode, max_iter=4, want_outputs=True)
        o = par.oracle_update(pts, small_scan["body"], small_scan["x_prior"], np.eye(npd) * 1e-4, mode, 4)
        assert r["trips"] == o["trips"] and r["searches"] == o["searches"] and r["n_eff"] == o["n_eff"]
        assert np.array_equal(r["eff"], o["eff"])
        assert np.abs(r["x"].as_array() - o["x"]).max() < 1e-12
    T.close()


def test_esti_plane_known_answers():
    rng = np.random.default_rng(0)
    for _ in range(200):
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        d = rng.uniform(0.5, 30.0)
        # five points exactly on the plane n.p + d = 0, spread over ~1 m
        u = np.cross(n, [1.0, 0.3, 0.2]); u /= np.linalg.norm(u); v = np.cross(n, u)
        ab = rng.uniform(-0.6, 0.6, (5, 2))
        P = (-d * n)[None, :] + ab[:, :1] * u + ab[:, 1:] * v
        ok, pabcd = orc.esti_plane(P.astype(np.float32))
        assert ok
        s = np.sign(np.dot(pabcd[:3], n))
        assert np.abs(pabcd[:3] * s - n).max() < 2e-3 and abs(pabcd[3] * s - d) < 2e-2
        # float64 least squares of the same system A x = -1
        x = np.linalg.lstsq(P.astype(np.float32).astype(np.float64), -np.ones(5), rcond=None)[0]
        nn = np.linalg.norm(x)
        assert np.abs(pabcd[:3] - x / nn).max() < 5e-3
        # square + centre, the centre lifted 1 m off the plane: no plane fits within the 0.1 m test
        sq = np.array([[-0.5, -0.5], [0.5, -0.5], [0.5, 0.5], [-0.5, 0.5], [0.0, 0.0]])
        P2 = (-d * n)[None, :] + sq[:, :1] * u + sq[:, 1:] * v
        P2[4] += 1.0 * n
        ok2, _ = orc.esti_plane(P2.astype(np.float32))
        assert not ok2


def _numpy_solve_v3_first_trip(P, HTH, HTz, R):
    """esekfom.hpp:1779-1812 on the first trip (dx = 0, Jacobians = I): two dense inverses in float64."""
    n = 23
    P_temp = np.linalg.inv(P / R)
    P_temp[:12, :12] += HTH
    P_inv = np.linalg.inv(P_temp)
    K_h = P_inv[:, :12] @ HTz
    return K_h


def test_solve_v3_first_trip_matches_numpy():
    rng = np.random.default_rng(1)
    A = rng.normal(size=(40, 12))
    HTH = A.T @ A * 50.0
    HTz = A.T @ rng.normal(size=40)
    B = rng.normal(size=(23, 23)); P = (B @ B.T) * 1e-5 + np.eye(23) * 1e-4
    x = orc.State.make(t=[1.0, 2.0, 3.0])
    xp = x.copy()
    r = orc.solve_v3(x, xp, P, HTH, HTz, 0.001, -1, 4, np.full(23, 0.001), 0)
    want = _numpy_solve_v3_first_trip(P, HTH, HTz, 0.001)
    assert np.abs(r["dx"] - want).max() < 1e-9 * max(1.0, np.abs(want).max())


def test_h_share_model_properties(small_world, small_scan):
    import flivo_parity as par
    cache = {}
    for mode in (orc.MODE_V3, orc.MODE_V1):
        o = par.oracle_trip(small_world["points"], small_scan["body"], small_scan["x_prior"], True, cache, mode)
        H = o["HTH"]
        assert np.array_equal(H, H.T)
        assert np.linalg.eigvalsh(H).min() > -1e-6 * np.abs(H).max()
        assert o["n_eff"] == int(o["eff"].sum()) and np.all(o["eff"] <= o["mask"])
        found5 = cache["cnt"] >= 5
        assert np.all(o["mask"] <= found5.astype(np.uint8))
        # V1 HPH/HPL is a column permutation of the V3 block scaled by 1/R
        HPH, HPL = orc.repack_v1(o["HTH"], o["HTz"], 0.001)
        assert np.allclose(HPH[:3, :3], H[3:6, 3:6] / 0.001) and np.allclose(HPH[3:6, 3:6], H[:3, :3] / 0.001)
        assert np.all(HPH[6:, :] == 0) and np.all(HPL[6:] == 0)


@pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built (needs /root/reference at build time)")
def test_edge_cases_against_live_reference_ikdtree():
    """Ragged and degenerate inputs on the reference tree run live: fewer than k points, a tight max_dist (0..k
    neighbours per query), coincident map points (distance ties: compared as distance lists), the half-open box of
    Delete_Point_Boxes, a query set of one, deleting everything."""
    rng = np.random.default_rng(11)
    # (1) fewer points than k
    few = rng.uniform(-1, 1, (3, 3)).astype(np.float32)
    T = orc.RefTree(0.5)
    T.build(few)
    q = rng.uniform(-2, 2, (50, 3)).astype(np.float32)
    rid, rsq, _, rcnt = T.knn(q, 5, np.inf)
    oid, osq, _, ocnt, _ = orc.knn_bruteforce(few, q, 5, np.inf)
    assert (rcnt == 3).all() and np.array_equal(rcnt, ocnt)
    assert np.array_equal(rid[:, :3], oid[:, :3]) and np.array_equal(rsq[:, :3], osq[:, :3])
    T.close()
    # (2) tight max_dist: ragged counts, and (3) coincident points
    pts = rng.uniform(-5, 5, (4000, 3)).astype(np.float32)
    pts[100:140] = pts[60:100]                                  # 40 exact duplicates
    T = orc.RefTree(0.5)
    T.build(pts)
    q = (pts[rng.integers(0, len(pts), 600)] + rng.normal(0, 0.05, (600, 3))).astype(np.float32)
    q[:40] = pts[60:100]                                        # queries on the duplicated points: distance 0 twice
    for md in (0.02, 0.2, np.inf):
        rid, rsq, _, rcnt = T.knn(q, 5, md)
        oid, osq, _, ocnt, ties = orc.knn_bruteforce(pts, q, 5, md)
        assert np.array_equal(rcnt, ocnt), md
        for i in range(len(q)):
            assert np.array_equal(rsq[i, :rcnt[i]], osq[i, :ocnt[i]]), (md, i)
        if md == 0.02:
            assert len(set(rcnt.tolist())) >= 4                 # really ragged
    one = T.knn(q[:1], 5, np.inf)
    assert one[3][0] == 5 and np.array_equal(one[1][0], orc.knn_bruteforce(pts, q[:1], 5, np.inf)[1][0])
    T.close()
    # (4) half-open delete box: min <= p < max per axis (ikd_Tree.cpp:648)
    grid = np.stack(np.meshgrid(np.arange(5), np.arange(5), np.arange(5), indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
    T, M = orc.RefTree(0.5), orc.OracleMap(0.5)
    T.build(grid)
    M.build(grid)
    box = np.array([[1, 1, 1, 3, 3, 3]], np.float32)            # faces on the lattice: planes 1 and 2 go, plane 3 stays
    assert T.delete_boxes(box) == M.delete_boxes(box) == 8
    rx, _ = T.flatten()
    ox, _ = M.live()
    key = lambda a: a[np.lexsort((a[:, 2], a[:, 1], a[:, 0]))]
    assert np.array_equal(key(rx), key(ox)) and len(ox) == 125 - 8
    assert not ((ox >= 1).all(1) & (ox < 3).all(1)).any() and ((ox == 3).any(1)).sum() > 0
    # (5) deleting everything, then searching
    allbox = np.array([[-10, -10, -10, 10, 10, 10]], np.float32)
    assert T.delete_boxes(allbox) == M.delete_boxes(allbox) == 117
    assert len(M.live()[0]) == 0 and (T.knn(grid[:5], 5, np.inf)[3] == 0).all()
    T.close()
