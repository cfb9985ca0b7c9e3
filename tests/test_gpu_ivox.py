"""-m gpu: iVox-compatible mode through the C ABI against the restated faster_lio::IVox
(oracle/ivox_oracle.cpp): AddPoints / GetClosestPoint, the h_share_model trip and the whole iterated update with
the iVox search, map_incremental's add rule, and the C++ drop-in class."""
import os
import struct
import subprocess

import numpy as np
import pytest

import flivo_parity as par
import flivo_scenes as scenes
from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu


def _dense_world(seed, n, spacing):
    """A city-block world whose points are `spacing` apart: several points per iVox voxel of 0.5 m."""
    return scenes.make_world(seed=seed, n_points=n, ds=spacing)


@pytest.mark.parametrize("nearby", [1, 7, 19, 27])
def test_addpoints_getclosestpoint_match_oracle(pkg, nearby):
    w = _dense_world(21, 300000, 0.15)
    pts = w["points"]
    rng = np.random.default_rng(5 + nearby)
    q = pts[rng.integers(0, len(pts), 5000)] + rng.normal(0, 0.04, (5000, 3)).astype(np.float32)
    q = np.vstack([q, rng.uniform(-30, 30, (500, 3)).astype(np.float32)]).astype(np.float32)
    v = orc.OracleIVox(0.5, nearby)
    with pkg.Flivo(downsample_size=0.5, initial_points=1 << 14, ivox_nearby=nearby) as F:   # small hint: tables must grow
        for part in np.array_split(pts, 3):
            F.AddPoints(part)
            v.add_points(part)
        assert F.size() == len(pts)
        gid, gsq, gxyz, gcnt = F.GetClosestPoint(q, 5, 5.0)
        oid, osq, oxyz, ocnt, ties = v.knn(q, 5, 5.0)
        assert ties == 0
        assert np.array_equal(gcnt, np.maximum(ocnt, 0))
        assert np.array_equal(gid, oid) and np.array_equal(gsq, osq) and np.array_equal(gxyz, oxyz)
        assert (ocnt == 5).sum() > 1000 and (ocnt < 0).sum() > 100
        if orc.have_ivox_ref():
            # the same against faster_lio::IVox ITSELF (include/ivox3d headers compiled unmodified, oracle/_ref/libivox_ref.so):
            # found / not found, count, nearest element and the neighbour set (the reference's order is nth_element's)
            R = orc.RefIVox(0.5, nearby)
            for part in np.array_split(pts, 3):
                R.add_points(part)
            rxyz, rcnt = R.knn(q, 5, 5.0)
            assert np.array_equal(gcnt, np.maximum(rcnt, 0)) and R.num_grids() == F.stats().map_cells
            has = gcnt > 0
            assert np.array_equal(gxyz[has, 0], rxyz[has, 0])
            for i in np.nonzero(has)[0]:
                a, b = gxyz[i, :gcnt[i]], rxyz[i, :gcnt[i]]
                assert np.array_equal(a[np.lexsort((a[:, 2], a[:, 1], a[:, 0]))], b[np.lexsort((b[:, 2], b[:, 1], b[:, 0]))]), i
        # voxels holding more than one bucket (8 slots) exist in this world: the overflow chains are exercised
        xyz, ids = F.flatten()
        key = np.round(xyz.astype(np.float32) * np.float32(1.0 / 0.5)).astype(np.int64)
        _, per_voxel = np.unique(key, axis=0, return_counts=True)
        assert per_voxel.max() > 8
        assert F.stats().map_cells == v.num_grids() == len(per_voxel)
        # a tighter range and a smaller k
        gid, gsq, _, gcnt = F.GetClosestPoint(q, 3, 0.2)
        oid, osq, _, ocnt, _ = v.knn(q, 3, 0.2)
        assert np.array_equal(gcnt, np.maximum(ocnt, 0)) and np.array_equal(gid, oid) and np.array_equal(gsq, osq)


@pytest.mark.parametrize("nearby", [1, 7, 19, 27])
def test_getclosestpoint_matches_reference_golden(pkg, nearby):
    """tests/golden/ivox_reference.npz: outputs of faster_lio::IVox itself (the reference's header compiled unmodified,
    tests/golden/make_golden.py) — found / not found, count, nearest element, neighbour set."""
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ivox_reference.npz"))
    with pkg.Flivo(downsample_size=float(gold["resolution"]), ivox_nearby=nearby) as F:
        def knn(q, k, max_range):
            _, _, xyz, cnt = F.GetClosestPoint(q, k, max_range)
            return xyz, cnt
        par.check_ivox_golden(gold, nearby, F.AddPoints, lambda: F.stats().map_cells, knn)


def test_ivox_mode_rejects_ikdtree_only_operations(pkg):
    pts = np.random.default_rng(3).uniform(-2, 2, (100, 3)).astype(np.float32)
    with pkg.Flivo(downsample_size=0.5, ivox_nearby=19) as F:
        F.AddPoints(pts)
        with pytest.raises(pkg.FlivoError):
            F.Add_Points(pts, True)           # voxel down-sampling is the ikd-Tree rule
        with pytest.raises(pkg.FlivoError):
            F.Delete_Points(pts[:3])
        assert F.size() == 100
        assert F.Delete_Point_Boxes(np.array([[-3, -3, -3, 0, 3, 3]], np.float32)) == int((pts[:, 0] < 0).sum())
    with pytest.raises(pkg.FlivoError):
        pkg.Flivo(downsample_size=0.5, ivox_nearby=5)


@pytest.mark.parametrize("mode,nearby,sort_scan", [(1, 19, False), (1, 19, True), (1, 7, True), (0, 27, True)])
def test_update_with_ivox_search_matches_oracle(pkg, mode, nearby, sort_scan):
    w = _dense_world(23, 80000, 0.2)
    scan = scenes.make_scan(w, seed=31, n_points=6000)
    v = orc.OracleIVox(0.5, nearby)
    v.add_points(w["points"])
    with pkg.Flivo(downsample_size=0.5, initial_points=len(w["points"]), sort_scan=sort_scan, ivox_nearby=nearby) as F:
        F.AddPoints(w["points"])
        F.scan_upload(scan["body"])
        g, o, cache = par.check_trip_ivox(pkg, F, v, scan["body"], scan["x_prior"], mode)
        assert g["n_eff"] > 2000
        xg, Pg, rep, oo = par.check_update_ivox(pkg, F, v, scan["body"], scan["x_prior"], mode)
        assert rep.searches >= 1
        t0 = np.linalg.norm(scan["x_prior"][9:12] - scan["x_true"][9:12])
        t1 = np.linalg.norm(xg.as_array()[9:12] - scan["x_true"][9:12])
        assert t1 < t0 or nearby < 19    # (a 7-voxel stencil on this sparse map is a poor estimator: parity is what is checked)


def test_empty_result_keeps_previous_neighbours(pkg):
    """GetClosestPoint returns before clearing closest_pt when no voxel of the stencil holds a point
    (ivox3d.h:168-170): a point that had neighbours in the first search keeps them when a later search finds none."""
    w = _dense_world(25, 30000, 0.2)
    scan = scenes.make_scan(w, seed=33, n_points=2000)
    v = orc.OracleIVox(0.5, 1)     # CENTER stencil: empty results are common
    v.add_points(w["points"])
    x0 = np.array(scan["x_true"], float)
    x1 = x0.copy()
    x1[9:12] += [0.4, -0.3, 0.25]   # a second state that moves most points into other voxels
    with pkg.Flivo(downsample_size=0.5, sort_scan=True, ivox_nearby=1) as F:
        F.AddPoints(w["points"])
        F.scan_upload(scan["body"])
        cache = {}
        par.oracle_trip(None, scan["body"], x0, True, cache, 1, ivox=v)
        g0 = F.h_share_model(pkg.State.from_array(x0), 1, True, want=("nbr_id",))
        had = cache["cnt"].copy()
        o1 = par.oracle_trip(None, scan["body"], x1, True, cache, 1, ivox=v)
        g1 = F.h_share_model(pkg.State.from_array(x1), 1, True, want=("nbr_id", "selected", "eff"))
        _, _, _, c1, _ = v.knn(orc.body_to_world(scan["body"], orc.State.from_array(x1)), 5, 5.0)
        kept = (c1 < 0) & (had > 0)
        assert kept.sum() > 10, "scene does not exercise the rule"
        want = np.where(np.arange(5)[None, :] < cache["cnt"][:, None], cache["ids"], -1)
        assert np.array_equal(g1["nbr_id"], want)
        assert np.array_equal(g1["nbr_id"][kept], g0["nbr_id"][kept])
        assert np.array_equal(g1["selected"], o1["mask"]) and np.array_equal(g1["eff"], o1["eff"])


@pytest.mark.parametrize("sort_scan", [False, True])
def test_map_incremental_ivox_branch(pkg, sort_scan):
    w = _dense_world(27, 60000, 0.2)
    fs = 0.4                                   # filter_size_map_min differs from the voxel resolution
    v = orc.OracleIVox(0.5, 19)
    v.add_points(w["points"])
    with pkg.Flivo(downsample_size=0.5, sort_scan=sort_scan, ivox_nearby=19) as F:
        F.AddPoints(w["points"])
        F.set_downsample_param(fs)
        base = len(w["points"])
        for k in range(3):
            scan = scenes.make_scan(w, seed=40 + k, n_points=5000)
            F.scan_upload(scan["body"])
            if k == 1:
                # a scan that was never searched: Nearest_Points empty -> every point is added
                x_post = np.array(scan["x_true"], float)
                near_xyz, near_cnt = np.zeros((5000, 5, 3), np.float32), np.zeros(5000, np.int32)
            else:
                xg, Pg, rep, o = par.check_update_ivox(pkg, F, v, scan["body"], scan["x_prior"], 1)
                x_post = xg.as_array()          # the state the caller hands to map_incremental
                near_xyz, near_cnt = o["xyz"], o["cnt"]
            wpts = orc.body_to_world(scan["body"], orc.State.from_array(x_post))
            flags = orc.ivox_map_filter(wpts, near_xyz, near_cnt, fs)
            add = np.vstack([wpts[flags == 1], wpts[flags == 2]])
            v.add_points(add)
            n_added = F.map_add_scan(pkg.State.from_array(x_post))
            assert n_added == len(add), "added %d vs oracle %d" % (n_added, len(add))
            if k != 1:
                assert 0 < (flags == 0).sum() and (flags == 1).sum() > 0
            base += len(add)
            assert F.size() == base
            xyz, ids = F.flatten()
            oxyz, oids = v.points()
            a, b = np.argsort(ids), np.argsort(oids)
            assert np.array_equal(ids[a], oids[b]) and np.array_equal(xyz[a], oxyz[b])


def test_ivox_dropin_class_cpp(pkg, tmp_path):
    import importlib
    exe = importlib.import_module("fast_livo_noted_b200.build").build_host_test()
    w = _dense_world(29, 30000, 0.2)
    pts = w["points"]
    rng = np.random.default_rng(9)
    q = np.vstack([pts[rng.integers(0, len(pts), 2000)] + rng.normal(0, 0.03, (2000, 3)).astype(np.float32),
                   rng.uniform(-40, 40, (100, 3)).astype(np.float32)]).astype(np.float32)
    rng.shuffle(q)
    fin, fout = os.path.join(tmp_path, "ivox_in.bin"), os.path.join(tmp_path, "ivox_out.bin")
    with open(fin, "wb") as f:
        f.write(struct.pack("4i", len(pts), len(q), 18, 0))
        f.write(struct.pack("d", 0.5))
        f.write(pts.astype(np.float32).tobytes())
        f.write(q.tobytes())
    r = subprocess.run([exe, "--ivox", fin, fout], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = open(fout, "rb").read()
    n_pts, n_grids, single_ok, _ = struct.unpack("4i", raw[:16])
    nq = len(q)
    ids = np.frombuffer(raw, np.int32, nq * 5, 16).reshape(nq, 5)
    cnt = np.frombuffer(raw, np.int32, nq, 16 + nq * 20)
    sqd = np.frombuffer(raw, np.float32, nq * 5, 16 + nq * 24).reshape(nq, 5)
    v = orc.OracleIVox(0.5, 19)
    v.add_points(pts)
    oid, osq, _, ocnt, ties = v.knn(q, 5, 5.0)
    assert ties == 0 and single_ok == 1
    assert n_pts == len(pts) and n_grids == v.num_grids()
    assert np.array_equal(cnt, np.maximum(ocnt, 0)) and np.array_equal(ids, oid) and np.array_equal(sqd, osq)


@pytest.mark.parametrize("ivox_nearby", [0, 19])
def test_gnss_block_folded_into_v1_update(pkg, ivox_nearby):
    """src/laser_mapping.cpp:612-617: the GNSS position block is added to HPH / HPL on every trip, on the device."""
    w = _dense_world(35, 80000, 0.2) if ivox_nearby else scenes.make_world(seed=35, n_points=60000, ds=0.5)
    scan = scenes.make_scan(w, seed=36, n_points=5000)
    lever, std = np.array([0.05, -0.1, 0.3]), np.array([0.05, 0.05, 0.1])
    xt = np.array(scan["x_true"], float)
    gpos = xt[9:12] - xt[0:9].reshape(3, 3) @ lever + np.array([0.04, -0.03, 0.02])   # antenna fix with a few cm of error
    v = None
    P0 = np.eye(18) * 1e-4
    with pkg.Flivo(downsample_size=0.5, initial_points=len(w["points"]), sort_scan=True, ivox_nearby=ivox_nearby) as F:
        if ivox_nearby:
            v = orc.OracleIVox(0.5, ivox_nearby)
            v.add_points(w["points"])
            F.AddPoints(w["points"])
        else:
            F.Build(w["points"])
        F.scan_upload(scan["body"])
        cfg = pkg.make_cfg(mode=pkg.MODE_V1, max_iter=4, laser_point_cov=0.001)
        x_plain, _, rep0 = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
        F.set_gnss(gpos, lever, std)
        xg, Pg, rep = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
        o = par.oracle_update(w["points"], scan["body"], scan["x_prior"], P0, orc.MODE_V1, 4, 0.001, None, False, ivox=v,
                              gnss=(gpos, lever, std))
        assert rep.trips == o["trips"] and rep.searches == o["searches"] and rep.n_eff == o["n_eff"]
        assert np.abs(np.array(rep.dx_last[:18]) - o["dx"][-1][:18]).max() <= par.DX_ATOL
        assert np.abs(xg.as_array() - o["x"]).max() <= 1e-6
        assert np.abs(Pg - o["P"]).max() / np.abs(o["P"]).max() <= 1e-6
        assert np.abs(xg.as_array() - x_plain.as_array()).max() > 1e-7, "the GNSS block changed nothing"
        # V3 ignores it (the block belongs to the compiled 18-DOF path); clearing restores the plain result
        F.set_gnss(None)
        x_again, _, _ = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
        assert np.array_equal(x_again.as_array(), x_plain.as_array())
        # an observation farther than 2 m from the state is dropped (src/GNSS_Processing.cpp:221-222)
        F.set_gnss(gpos + np.array([5.0, 0, 0]), lever, std)
        x_far, _, _ = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
        assert np.array_equal(x_far.as_array(), x_plain.as_array())
