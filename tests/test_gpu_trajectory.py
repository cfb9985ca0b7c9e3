"""-m gpu: a trajectory in the style of BASELINE config 2 — every frame: iterated update against the current map, then
map_incremental (Add_Points with down-sampling) — against the oracle run on the same growing map.  Both flavours.
The neighbour sets, masks and trip policy must agree at every frame, otherwise the maps diverge and later frames fail."""
import numpy as np
import pytest

import flivo_parity as par
import flivo_scenes as scenes
from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", [0, 1])
def test_trajectory_update_then_add_points(pkg, mode):
    world = scenes.make_world(seed=21, n_points=30000, ds=0.5)
    poses = scenes.street_positions(world, 6)
    npd = 18 if mode == 1 else 23
    M = orc.OracleMap(world["ds"])
    M.build(world["points"])
    cur = world["points"].copy()
    with pkg.Flivo(downsample_size=world["ds"], initial_points=len(cur)) as F:
        F.Build(cur)
        for k, pos in enumerate(poses):
            scan = scenes.make_scan(world, seed=300 + k, n_points=2500, pos=pos)
            P0 = np.eye(npd) * 1e-4
            o = par.oracle_update(cur, scan["body"], scan["x_prior"], P0, mode, 4, v1_bound=pkg.v1_search_bound_sq(world["ds"]))
            F.scan_upload(scan["body"])
            x, P, rep = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, mode=mode, max_iter=4)
            assert (rep.trips, rep.searches, rep.n_eff) == (o["trips"], o["searches"], o["n_eff"]), "frame %d" % k
            assert np.abs(x.as_array() - o["x"]).max() < 1e-6
            assert np.abs(P - o["P"]).max() <= 1e-6 * np.abs(o["P"]).max()
            # map_incremental with the ORACLE's posterior on both sides, so that one ulp of state difference cannot
            # move a point across a voxel boundary and hide or fake a map mismatch
            st = orc.State.from_array(o["x"])
            added = M.add_points(orc.body_to_world(scan["body"], st), True)
            assert F.map_add_scan(pkg.State.from_array(o["x"])) == added
            cur, ids = M.live()
            gx, gid = F.flatten()
            og, oo = np.argsort(gid), np.argsort(ids)
            assert np.array_equal(gid[og], ids[oo]) and np.array_equal(gx[og], cur[oo]), "map differs after frame %d" % k
