"""-m gpu: the super-block occupancy table (csrc/flivo_dev.cuh: sup_keys / sup_mask) that wide search cubes go through.
Queries that look into unmapped space from 5 .. 30 cells away make the warp-cooperative search walk cubes of more than 64
blocks, where only the blocks the table reports are probed: a block the table misses is a neighbour lost.  The map is grown
in many small batches from a tiny initial capacity, so blocks are created by Add_Points, by Build, and carried through
several directory rehashes; results are compared with the oracle's brute force (float32 distances as the kernels form them,
ties broken by insertion id)."""
import numpy as np
import pytest

import flivo_parity as par

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("ds,ratio", [(0.5, 2), (0.2, 2)])
def test_wide_cubes_after_growth_and_rehash(pkg, ds, ratio):
    rng = np.random.default_rng(77)
    cell = ds * ratio
    # islands of points far apart (several super-blocks each, many empty ones between), added batch by batch
    centres = rng.uniform(-40, 40, (30, 3)) * cell
    with pkg.Flivo(downsample_size=ds, cell_ratio=ratio, initial_points=256, sort_scan=True) as F:
        cap0 = F.stats().dir_capacity
        first = (centres[0] + rng.uniform(-3, 3, (400, 3)) * cell).astype(np.float32)
        F.Build(first)
        for c in centres[1:15]:
            F.Add_Points((c + rng.uniform(-6, 6, (1500, 3)) * cell).astype(np.float32), downsample_on=True)
        # one batch large enough that its conservative sizing outgrows the 65536-slot directory: the blocks made so far are
        # rehashed (k_rehash marks them in the new super-block table), the rest is added afterwards
        big = np.concatenate([c + rng.uniform(-6, 6, (2700, 3)) * cell for c in centres[15:30]]).astype(np.float32)
        assert len(big) > 30000
        F.Add_Points(big, downsample_on=True)
        assert F.stats().dir_capacity > cap0, "the directory never grew: the test does not cover the rehash"
        mxyz, mid = F.flatten()
        # queries 5 .. 30 cells away from an island, looking at it across empty space, and a few inside
        q = []
        for c in centres[::3]:
            for r in (5, 9, 14, 22, 30):
                v = rng.normal(size=3); v /= np.linalg.norm(v)
                q.append(c + v * (6 + r) * cell)
            q.append(c + rng.uniform(-1, 1, 3) * cell)
        q = np.asarray(q, np.float32)
        max_dist = (31.5 * cell) ** 2                   # KD_TREE::Nearest_Search compares SQUARED distances with max_dist
        gid, gsq, gxyz, gcnt = F.Nearest_Search(q, 5, max_dist)
        oid, osq, oxyz, ocnt, ties = par.oracle_knn(mxyz, q, 5, max_dist, ids=mid)
        assert ties == 0
        assert np.array_equal(gcnt, ocnt)
        assert np.array_equal(gid, oid)
        assert np.array_equal(gsq, osq)
        assert (gcnt > 0).sum() >= len(q) // 2          # the far queries did reach an island
