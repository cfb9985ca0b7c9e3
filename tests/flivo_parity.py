"""Parity checks of the CUDA path (through the C ABI) against the CPU oracle.  Used by tests/ (-m gpu) and by
__graft_entry__.smoke().  The oracle is the checker here, never the thing measured.

Tolerances (BASELINE.json north_star): neighbour index sets and valid-point masks bit-exact; H^T H / H^T z
within 1e-5 relative per entry (relative to max(|entry|, 1e-12 * ||HTH||_F)); per-trip state
increment within 1e-6 absolute per component.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from oracle import pyoracle as orc  # noqa: E402

HTH_RTOL = 1e-5
DX_ATOL = 1e-6
GATE_SQ = 5.0   # pointSearchSqDis[4] > 5 (src/laser_mapping.cpp:518): the fused path's search bound


def oracle_knn(map_xyz, queries, k=5, max_dist=np.inf, ids=None, alive=None):
    return orc.knn_bruteforce(map_xyz, queries, k, max_dist, ids, alive)


def oracle_trip(map_xyz, body, x, searched, cache, mode, mask_prev=None, map_ids=None, ivox=None, v1_bound=None):
    """One h_share_model trip on the CPU.  cache: dict carrying Nearest_Points between trips.
    ivox: an orc.OracleIVox — the search is then IVox::GetClosestPoint (src/laser_mapping.cpp:520) and the
    far-neighbour gate of the ikd-Tree branch (:518) does not exist."""
    st = orc.State.from_array(x)
    if searched and ivox is not None:
        wpts = orc.body_to_world(body, st)
        oid, osq, oxyz, ocnt, ties = ivox.knn(wpts, 5, 5.0)
        none = ocnt < 0                       # no candidate: the list keeps its previous content (ivox3d.h:168-170)
        if "ids" in cache:
            for new, key in ((oid, "ids"), (osq, "sqd"), (oxyz, "xyz"), (ocnt, "cnt")):
                new[none] = cache[key][none]
        else:
            ocnt[none] = 0
        cache.update(ids=oid, sqd=osq, xyz=oxyz, cnt=ocnt, ties=ties + cache.get("ties", 0))
    elif searched:
        wpts = orc.body_to_world(body, st)
        # V1 keeps neighbours up to FLIVO_V1_RANGE_CELLS cells only (stated deviation, include/flivo.h): same bound here
        oid, osq, oxyz, ocnt, ties = oracle_knn(map_xyz, wpts, 5, np.inf if (mode != orc.MODE_V1 or v1_bound is None) else v1_bound, ids=map_ids)
        cache.update(ids=oid, sqd=osq, xyz=oxyz, cnt=ocnt, ties=ties)
    if ivox is not None:
        d5 = np.zeros(len(body), np.float32)
    else:
        d5 = np.where(cache["cnt"] >= 5, cache["sqd"][:, 4], np.inf).astype(np.float32)
    out = orc.h_share_model(body, st, cache["xyz"], cache["cnt"], d5, searched, mode, mask_prev)
    return out


def oracle_update(map_xyz, body, x_prior, P, mode, max_iter, laser_point_cov=0.001, limit=None, force_search=False,
                  ivox=None, gnss=None, v1_bound=None):
    """The whole iterated update with the restated reference loop (V3: esekfom.hpp:1619-1928; V1:
    src/laser_mapping.cpp:165-242).  Returns dict(x, P, trips, searches, dx (per trip), masks, eff, ids)."""
    limit = np.full(23, 0.001) if limit is None else np.asarray(limit, float)
    x = orc.State.from_array(x_prior)
    x_prop = orc.State.from_array(x_prior)
    P = np.array(P, dtype=np.float64, copy=True)
    P_prop = P.copy()
    cache = {}
    mask = np.ones(len(body), np.uint8)
    search, t, rematch, stopped = 1, 0, 0, 0
    trips = searches = 0
    dxs, trip_out = [], []
    it = -1
    out = None
    while it < max_iter and not stopped:
        do_search = bool(search or force_search)
        out = oracle_trip(map_xyz, body, x.as_array(), do_search, cache, mode, mask, ivox=ivox, v1_bound=v1_bound)
        mask = out["mask"]
        searches += int(do_search)
        if mode == orc.MODE_V1:
            HPH, HPL = orc.repack_v1(out["HTH"], out["HTz"], laser_point_cov)
            if gnss is not None:                      # src/laser_mapping.cpp:612-617
                gH, gL = orc.gnss_block(x, *gnss)
                HPH = HPH.copy(); HPL = HPL.copy()
                HPH[:6, :6] += gH
                HPL[:6] += gL
            r = orc.solve_v1(x, x_prop, P, HPH, HPL, it, max_iter, rematch)
            rematch, search, stopped = r["rematch"], r["search_next"], r["stopped"]
            dxs.append(r["solution"].copy())
            if stopped:
                P = r["cov"]
        else:
            r = orc.solve_v3(x, x_prop, P_prop, out["HTH"], out["HTz"], laser_point_cov, it, max_iter, limit, t)
            t, search, stopped = r["t"], r["converge"], r["stopped"]
            dxs.append(r["dx"].copy())
            if stopped:
                P = r["P"]
        trip_out.append(dict(searched=do_search, n_eff=out["n_eff"], HTH=out["HTH"].copy(), HTz=out["HTz"].copy(),
                             x_after=x.as_array()))
        trips += 1
        it += 1
    return dict(x=x.as_array(), P=P, trips=trips, searches=searches, dx=dxs, mask=mask, eff=out["eff"],
                ids=cache["ids"], cnt=cache["cnt"], sqd=cache["sqd"], xyz=cache["xyz"], ties=cache["ties"], trip_out=trip_out,
                n_eff=out["n_eff"], borderline=out["borderline"])


def rel_err_block(got, want):
    want = np.asarray(want, float)
    got = np.asarray(got, float)
    floor = 1e-12 * np.linalg.norm(want)
    return np.max(np.abs(got - want) / np.maximum(np.abs(want), max(floor, 1e-300)))


def compare_neighbours(gpu_ids, orc_ids, orc_sqd, orc_cnt, bound=GATE_SQ):
    """GPU ids (searched with `bound` as max squared distance) vs the unbounded oracle list: row by row the GPU
    list must equal the oracle's entries with d <= bound (a prefix, both ascending), -1 padded.  Returns the
    number of mismatching rows."""
    want = np.where((orc_sqd <= np.float32(bound)) & (np.arange(5)[None, :] < orc_cnt[:, None]), orc_ids, -1)
    return int(np.any(want != gpu_ids, axis=1).sum()), want


def check_trip(pkg, F, map_xyz, body, x, mode, verbose=False):
    """Single searched trip: neighbour ids, masks, planes, world points, H^T H, H^T z."""
    st = pkg.State.from_array(x)
    g = F.h_share_model(st, mode, True, want=("selected", "eff", "nbr_id", "nbr_sqd", "plane", "world"))
    cache = {}
    o = oracle_trip(map_xyz, body, x, True, cache, mode, v1_bound=pkg.v1_search_bound_sq(F.ds, F.cell_ratio))
    assert cache["ties"] == 0, "synthetic data has a float32 tie straddling rank 5; re-seed"
    # V3 searches with the far-neighbour gate as its bound (equivalent there); V1 searches unbounded like the reference
    bad, want_ids = compare_neighbours(g["nbr_id"], cache["ids"], cache["sqd"], cache["cnt"],
                                       pkg.v1_search_bound_sq(F.ds, F.cell_ratio) if mode == orc.MODE_V1 else GATE_SQ)
    assert bad == 0, "%d / %d neighbour rows differ" % (bad, len(body))
    assert np.array_equal(g["world"], o["world"]), "feats_down_world differs"
    sqd_want = np.where(want_ids >= 0, cache["sqd"], np.inf).astype(np.float32)
    assert np.array_equal(g["nbr_sqd"], sqd_want), "neighbour squared distances differ"
    assert np.array_equal(g["selected"], o["mask"]), "point_selected_surf differs at %d points" % int((g["selected"] != o["mask"]).sum())
    assert np.array_equal(g["eff"], o["eff"]), "effective-row mask differs"
    assert g["n_eff"] == o["n_eff"]
    assert np.array_equal(g["plane"], o["normvec"]), "normvec differs (max %g)" % np.abs(g["plane"] - o["normvec"]).max()
    e1, e2 = rel_err_block(g["HTH"], o["HTH"]), rel_err_block(g["HTz"], o["HTz"])
    assert e1 <= HTH_RTOL and e2 <= HTH_RTOL, "HTH / HTz relative error %g / %g" % (e1, e2)
    assert abs(g["res_sum"] - o["res_sum"]) <= 1e-9 * max(1.0, abs(o["res_sum"]))
    if verbose:
        print("trip: n=%d n_eff=%d selected=%d borderline=%d HTH rel err %.2e HTz rel err %.2e" %
              (len(body), g["n_eff"], int(g["selected"].sum()), o["borderline"], e1, e2))
    return g, o


def check_update_against_oracle(pkg, world, scan, mode=0, max_iter=4, force_search=False, verbose=False,
                                sort_scan=False, cell_ratio=2):
    """Build the map on the GPU, run one checked trip and the whole iterated update, compare with the oracle."""
    map_xyz, body = world["points"], scan["body"]
    npd = 18 if mode == pkg.MODE_V1 else 23
    P0 = np.eye(npd) * 1e-4
    with pkg.Flivo(downsample_size=world["ds"], cell_ratio=cell_ratio, initial_points=len(map_xyz), sort_scan=sort_scan) as F:
        F.Build(map_xyz)
        assert F.size() == len(map_xyz)
        F.scan_upload(body)
        check_trip(pkg, F, map_xyz, body, scan["x_prior"], mode, verbose)
        cfg = pkg.make_cfg(mode=mode, max_iter=max_iter, laser_point_cov=0.001, force_search=force_search)
        xg, Pg, rep = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
        o = oracle_update(map_xyz, body, scan["x_prior"], P0, mode, max_iter, 0.001, None, force_search,
                          v1_bound=pkg.v1_search_bound_sq(world["ds"], cell_ratio))
        assert rep.trips == o["trips"], "trips %d vs oracle %d" % (rep.trips, o["trips"])
        assert rep.searches == o["searches"], "searches %d vs oracle %d" % (rep.searches, o["searches"])
        assert rep.n_eff == o["n_eff"], "n_eff %d vs oracle %d" % (rep.n_eff, o["n_eff"])
        nd = 18 if mode == pkg.MODE_V1 else 23
        dx_err = np.abs(np.array(rep.dx_last[:nd]) - o["dx"][-1][:nd]).max()
        assert dx_err <= DX_ATOL, "last state increment differs by %g" % dx_err
        x_err = np.abs(xg.as_array() - o["x"]).max()
        assert x_err <= 1e-6, "posterior state differs by %g" % x_err
        p_err = np.abs(Pg - o["P"]).max() / np.abs(o["P"]).max()
        assert p_err <= 1e-6, "posterior covariance differs by %g (relative to max entry)" % p_err
        # the posterior must have moved toward the truth
        t_err0 = np.linalg.norm(scan["x_prior"][9:12] - scan["x_true"][9:12])
        t_err1 = np.linalg.norm(xg.as_array()[9:12] - scan["x_true"][9:12])
        if verbose:
            print("update: trips=%d searches=%d n_eff=%d dx_err=%.2e x_err=%.2e P_err=%.2e pos err %.4f -> %.4f m gpu %.3f ms" %
                  (rep.trips, rep.searches, rep.n_eff, dx_err, x_err, p_err, t_err0, t_err1, rep.gpu_ms))
        return dict(rep=rep, x=xg, P=Pg, oracle=o, pos_err=(t_err0, t_err1))


# ---- iVox mode --------------------------------------------------------------------------------
def check_trip_ivox(pkg, F, ivox, body, x, mode, verbose=False):
    """Single searched trip in iVox mode against the restated IVox (oracle/ivox_oracle.cpp)."""
    st = pkg.State.from_array(x)
    g = F.h_share_model(st, mode, True, want=("selected", "eff", "nbr_id", "nbr_sqd", "plane", "world"))
    cache = {}
    o = oracle_trip(None, body, x, True, cache, mode, ivox=ivox)
    assert cache["ties"] == 0, "synthetic data has a float32 distance tie across rank 5; re-seed"
    want_ids = np.where(np.arange(5)[None, :] < cache["cnt"][:, None], cache["ids"], -1)
    bad = int(np.any(want_ids != g["nbr_id"], axis=1).sum())
    assert bad == 0, "%d / %d neighbour rows differ" % (bad, len(body))
    assert np.array_equal(g["world"], o["world"]), "feats_down_world differs"
    assert np.array_equal(g["nbr_sqd"], np.where(want_ids >= 0, cache["sqd"], np.inf).astype(np.float32))
    assert np.array_equal(g["selected"], o["mask"]), "point_selected_surf differs at %d points" % int((g["selected"] != o["mask"]).sum())
    assert np.array_equal(g["eff"], o["eff"]), "effective-row mask differs"
    assert g["n_eff"] == o["n_eff"]
    assert np.array_equal(g["plane"], o["normvec"]), "normvec differs"
    e1, e2 = rel_err_block(g["HTH"], o["HTH"]), rel_err_block(g["HTz"], o["HTz"])
    assert e1 <= HTH_RTOL and e2 <= HTH_RTOL, "HTH / HTz relative error %g / %g" % (e1, e2)
    if verbose:
        print("ivox trip: n=%d n_eff=%d five-neighbour rows=%d HTH rel err %.2e" % (len(body), g["n_eff"], int((cache["cnt"] >= 5).sum()), e1))
    return g, o, cache


def check_update_ivox(pkg, F, ivox, body, x_prior, mode, max_iter=4, verbose=False):
    """Whole iterated update in iVox mode against the restated loop."""
    npd = 18 if mode == pkg.MODE_V1 else 23
    P0 = np.eye(npd) * 1e-4
    cfg = pkg.make_cfg(mode=mode, max_iter=max_iter, laser_point_cov=0.001)
    xg, Pg, rep = F.iekf_update(pkg.State.from_array(x_prior), P0, cfg)
    o = oracle_update(None, body, x_prior, P0, mode, max_iter, 0.001, None, False, ivox=ivox)
    assert o["ties"] == 0
    assert rep.trips == o["trips"] and rep.searches == o["searches"], "trips/searches %d/%d vs oracle %d/%d" % (rep.trips, rep.searches, o["trips"], o["searches"])
    assert rep.n_eff == o["n_eff"], "n_eff %d vs oracle %d" % (rep.n_eff, o["n_eff"])
    dx_err = np.abs(np.array(rep.dx_last[:npd]) - o["dx"][-1][:npd]).max()
    assert dx_err <= DX_ATOL, "last state increment differs by %g" % dx_err
    x_err = np.abs(xg.as_array() - o["x"]).max()
    assert x_err <= 1e-6, "posterior state differs by %g" % x_err
    p_err = np.abs(Pg - o["P"]).max() / np.abs(o["P"]).max()
    assert p_err <= 1e-6
    if verbose:
        print("ivox update: trips=%d searches=%d n_eff=%d dx_err=%.2e x_err=%.2e" % (rep.trips, rep.searches, rep.n_eff, dx_err, x_err))
    return xg, Pg, rep, o


# ---- iVox golden vectors (tests/golden/ivox_reference.npz: outputs of the reference's own compiled header) -----------
def lexsorted_rows(xyz, cnt):
    """Every row's first cnt points in lexicographic order, the rest zero (same helper as tests/golden/make_golden.py)."""
    out = np.zeros_like(xyz)
    for i in range(len(cnt)):
        c = max(int(cnt[i]), 0)
        r = xyz[i, :c]
        out[i, :c] = r[np.lexsort((r[:, 2], r[:, 1], r[:, 0]))]
    return out


def check_ivox_golden(gold, nearby, add_points, num_grids, knn):
    """add_points(xyz), num_grids() -> int, knn(q, k, max_range) -> (xyz [n][k][3], count with <= 0 meaning none) of a map
    created for stencil `nearby` at gold['resolution'].  Found / not found, count, nearest element and neighbour set must
    be the reference's."""
    for part in np.array_split(gold["map"], 3):
        add_points(part)
    assert num_grids() == int(gold["grids_%d" % nearby])
    q = gold["queries"]
    for tag, k, max_range in (("k5", 5, 5.0), ("k3r02", 3, 0.2)):
        xyz, cnt = knn(q, k, max_range)
        cnt = np.maximum(np.asarray(cnt), 0)
        gcnt = np.maximum(gold["cnt_%d_%s" % (nearby, tag)], 0)
        assert np.array_equal(cnt, gcnt), (nearby, tag)
        has = cnt > 0
        assert has.sum() > 300 and (~has).sum() > 20
        assert np.array_equal(xyz[has, 0], gold["first_%d_%s" % (nearby, tag)][has]), (nearby, tag)
        assert np.array_equal(lexsorted_rows(xyz, cnt), gold["set_%d_%s" % (nearby, tag)]), (nearby, tag)


# ---- sharded update with every rank in THIS process ---------------------------------------------
def sharded_update_in_process(pkg, world, scan, nranks=2, mode=0, max_iter=4, device=0, verbose=False, delay_rank=None,
                              delay_s=0.0, timeout_s=None, ctxs=None, sort_scan=True):
    """The scan is cut into `nranks` contiguous blocks (sharding.shard_range); every rank = one context with the whole map
    and its block, driven by its own host thread; the per-trip exchange goes through the ranks' exchange buffers
    (flivo_peer_attach_local: the peer phase of the solve, exactly as between GPUs, minus the NVLink hop).  Checks: every
    rank ends with the bit-identical posterior, and that posterior is the oracle's for the WHOLE scan (x <= 1e-6, same
    trips / searches / n_eff).  Returns dict(x, P, reps, oracle, errors)."""
    import threading
    from importlib import import_module
    shard_range = import_module(pkg.__name__ + ".sharding").shard_range
    map_xyz, body = world["points"], scan["body"]
    npd = 18 if mode == pkg.MODE_V1 else 23
    P0 = np.eye(npd) * 1e-4
    cfg = pkg.make_cfg(mode=mode, max_iter=max_iter, laser_point_cov=0.001)
    own = ctxs is None
    if own:
        ctxs = [pkg.Flivo(device=device, downsample_size=world["ds"], initial_points=len(map_xyz), sort_scan=sort_scan) for _ in range(nranks)]
        for F in ctxs:
            F.Build(map_xyz)
        bases = [F.peer_local_base() for F in ctxs]
        for r, F in enumerate(ctxs):
            F.peer_attach_local(bases, r, nranks)
    if timeout_s is not None:
        for F in ctxs:
            F.peer_set_timeout(timeout_s)
    for r, F in enumerate(ctxs):
        lo, hi = shard_range(len(body), r, nranks)
        F.scan_upload(np.ascontiguousarray(body[lo:hi]))
    out, errs = [None] * nranks, [None] * nranks

    def run(r):
        import time
        try:
            if delay_rank == r:
                time.sleep(delay_s)
            out[r] = ctxs[r].iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
        except pkg.FlivoError as e:
            errs[r] = e

    th = [threading.Thread(target=run, args=(r,)) for r in range(nranks)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    res = dict(ctxs=ctxs, errors=errs, out=out)
    if any(errs):
        return res
    x0, Pm0, rep0 = out[0]
    for r in range(1, nranks):
        assert np.array_equal(out[r][0].as_array(), x0.as_array()), "rank %d posterior differs from rank 0 (must be bit-identical)" % r
        assert np.array_equal(out[r][1], Pm0)
    o = oracle_update(map_xyz, body, scan["x_prior"], P0, mode, max_iter, 0.001, None, False, v1_bound=pkg.v1_search_bound_sq(world["ds"]))
    assert rep0.trips == o["trips"] and rep0.searches == o["searches"], "trips/searches %d/%d vs oracle %d/%d" % (rep0.trips, rep0.searches, o["trips"], o["searches"])
    assert rep0.n_eff == o["n_eff"], "n_eff (all ranks) %d vs oracle %d" % (rep0.n_eff, o["n_eff"])
    x_err = np.abs(x0.as_array() - o["x"]).max()
    assert x_err <= 1e-6, "sharded posterior differs from the oracle's by %g" % x_err
    p_err = np.abs(Pm0 - o["P"]).max() / np.abs(o["P"]).max()
    assert p_err <= 1e-6
    if verbose:
        print("sharded in-process: ranks=%d trips=%d n_eff=%d x_err=%.2e P_err=%.2e" % (nranks, rep0.trips, rep0.n_eff, x_err, p_err))
    res.update(x=x0, P=Pm0, rep=rep0, oracle=o, x_err=x_err)
    if own:
        for F in ctxs:
            F.close()
    return res
