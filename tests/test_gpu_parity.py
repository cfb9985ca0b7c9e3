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


def _fringe_scene():
    """A map with a dense planar patch near the sensor and a SPARSE ground (2.5 m lattice, z ~ 0) further out: scan points
    landing on the sparse part find their five nearest neighbours beyond the far-neighbour gate (squared distance > 5,
    src/laser_mapping.cpp:518) and yet coplanar within esti_plane's 0.1 m."""
    rng = np.random.default_rng(77)
    gx, gy = np.meshgrid(np.arange(-12, 12, 0.5), np.arange(-12, 12, 0.5))
    dense = np.stack([gx.ravel(), gy.ravel(), rng.normal(0, 0.005, gx.size)], 1)
    sx, sy = np.meshgrid(np.arange(-60, 60, 2.5), np.arange(-60, 60, 2.5))
    keep = (np.abs(sx) > 14) | (np.abs(sy) > 14)
    sparse = np.stack([sx[keep] + 0.3, sy[keep] + 0.2, rng.normal(0, 0.005, keep.sum())], 1)
    wall = np.stack([np.full(1600, 11.9), rng.uniform(-10, 10, 1600), rng.uniform(0.3, 6, 1600)], 1)   # gives the update its yaw / x
    wall2 = np.stack([rng.uniform(-10, 10, 1600), np.full(1600, -11.9), rng.uniform(0.3, 6, 1600)], 1)
    pts = scenes.voxel_downsample(np.concatenate([dense, sparse, wall, wall2]).astype(np.float32), 0.5)
    world = dict(points=pts, ds=0.5)
    # scan: ground hits on a fan of rays from a sensor 2 m above the origin + wall hits, in the lidar frame of x_true
    x_true = np.zeros(36); x_true[0:9] = np.eye(3).ravel(); x_true[9:12] = [0.2, -0.1, 2.0]
    x_true[12:21] = np.eye(3).ravel(); x_true[21:24] = scenes.T_LI; x_true[33:36] = [0, 0, -9.809]
    r = np.concatenate([rng.uniform(2, 11, 2500), rng.uniform(16, 55, 700)])
    a = rng.uniform(0, 2 * np.pi, r.size)
    ground = np.stack([r * np.cos(a), r * np.sin(a), rng.normal(0, 0.01, r.size)], 1)
    w1 = np.stack([np.full(600, 11.9) + rng.normal(0, 0.01, 600), rng.uniform(-9, 9, 600), rng.uniform(0.5, 5.5, 600)], 1)
    w2 = np.stack([rng.uniform(-9, 9, 600), np.full(600, -11.9) + rng.normal(0, 0.01, 600), rng.uniform(0.5, 5.5, 600)], 1)
    wpts = np.concatenate([ground, w1, w2])
    body = (wpts - x_true[9:12] - x_true[21:24]).astype(np.float32)          # R = R_LI = I
    x_prior = x_true.copy()
    x_prior[9:12] += [0.03, -0.02, 0.04]
    x_prior[0:9] = scenes.exp_so3(np.array([0.003, -0.002, 0.004])).ravel()
    return world, dict(body=body, x_prior=x_prior, x_true=x_true)


def test_v1_reuse_trips_keep_points_beyond_the_far_neighbour_gate(pkg):
    """in V1 point_selected_surf is a local vector<bool>(n, true) per call (src/laser_mapping.cpp:490)
    and the `> 5` gate acts on search trips only (:518), so a point whose five neighbours lie beyond the gate re-enters on
    re-use trips when its plane fits.  The GPU path must carry those five neighbours (unbounded search in V1)."""
    from oracle import pyoracle as orc
    world, scan = _fringe_scene()
    map_xyz, body, x0 = world["points"], scan["body"], scan["x_prior"]
    with pkg.Flivo(downsample_size=world["ds"], initial_points=len(map_xyz)) as F:
        F.Build(map_xyz)
        F.scan_upload(body)
        st = pkg.State.from_array(x0)
        cache = {}
        o_s = par.oracle_trip(map_xyz, body, x0, True, cache, orc.MODE_V1)
        assert cache["ties"] == 0
        fringe = (cache["cnt"] >= 5) & (cache["sqd"][:, 4] > 5.0)
        assert fringe.sum() > 200, "scene has no fringe points"
        g_s = F.h_share_model(st, pkg.MODE_V1, True, want=("selected", "eff", "nbr_id"))
        bound = pkg.v1_search_bound_sq(world["ds"])      # V1 keeps neighbours up to 24 cells (12 m here): beyond the gate
        assert bound > 100.0 and (cache["sqd"][fringe][:, 4] < bound).all()
        assert np.array_equal(g_s["nbr_id"], np.where(np.arange(5)[None, :] < cache["cnt"][:, None], cache["ids"], -1))
        assert np.array_equal(g_s["selected"], o_s["mask"]) and g_s["n_eff"] == o_s["n_eff"]
        assert not o_s["eff"][fringe].any()                      # search trip: gated out
        o_r = par.oracle_trip(map_xyz, body, x0, False, cache, orc.MODE_V1, o_s["mask"])
        g_r = F.h_share_model(st, pkg.MODE_V1, False, want=("selected", "eff"))
        assert o_r["eff"][fringe].sum() > 100                    # re-use trip: they are back (reference semantics)
        assert np.array_equal(g_r["eff"], o_r["eff"]) and np.array_equal(g_r["selected"], o_r["mask"])
        assert g_r["n_eff"] == o_r["n_eff"] and g_r["n_eff"] > g_s["n_eff"]
        assert par.rel_err_block(g_r["HTH"], o_r["HTH"]) <= par.HTH_RTOL
        # V3 (persistent flags): the same points stay out on the re-use trip
        F.scan_upload(body)
        g3 = F.h_share_model(st, pkg.MODE_V3, True, want=("selected",))
        g3r = F.h_share_model(st, pkg.MODE_V3, False, want=("eff",))
        assert not g3r["eff"][fringe].any()
    par.check_update_against_oracle(pkg, world, scan, mode=pkg.MODE_V1, max_iter=4, verbose=True)
    par.check_update_against_oracle(pkg, world, scan, mode=pkg.MODE_V3, max_iter=4, verbose=True)
