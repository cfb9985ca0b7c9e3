#!/usr/bin/env python
"""Workload of the compute-sanitizer record (scripts/gpu_sanitize.sh runs it under memcheck / racecheck /
synccheck / initcheck): every kernel family of libflivo_b200.so once at smoke size, checked against the oracle as it goes.

  update   : the iterated update in both modes (V3, V1) and with force_search — k_search (tile phase, far queue, plane
             fit), k_fused (DMMA Gram, group fold with the last-arriver ticket), the solve, scan staging
  map      : the ikd-Tree golden sequence (Build, Add_Points with/without down-sampling, box delete, kNN)
  peer     : a 2-rank sharded update on ONE GPU (two contexts of this process exchanging their normal-equation records
             through each other's exchange buffers: the peer flags / release-acquire protocol of the solve)
Sizes are small because racecheck slows kernels down by two orders of magnitude.
"""
import os
import sys
import threading

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import __graft_entry__ as ge  # noqa: E402


def run_update(pkg, n_map=60000, n_scan=4000):
    import flivo_scenes as scenes
    from flivo_parity import check_update_against_oracle
    world = scenes.make_world(seed=7, n_points=n_map, ds=0.5)
    scan = scenes.make_scan(world, seed=8, n_points=n_scan)
    check_update_against_oracle(pkg, world, scan, mode=pkg.MODE_V3, max_iter=4, verbose=True)
    check_update_against_oracle(pkg, world, scan, mode=pkg.MODE_V1, max_iter=4, verbose=True)
    check_update_against_oracle(pkg, world, scan, mode=pkg.MODE_V3, max_iter=3, force_search=True, verbose=True)
    print("update ok")


def run_map(pkg):
    g = np.load(os.path.join(ROOT, "tests", "golden", "map_ops_ikdtree.npz"))
    with pkg.Flivo(downsample_size=float(g["ds"]), cell_ratio=2, initial_points=1 << 14) as F:
        F.Build(g["base"])
        for b in range(3):
            assert F.Add_Points(g["batch%d" % b], True) == int(g["ret%d" % b])
        assert F.Add_Points(g["batch_nods"], False) == int(g["ret_nods"])
        assert F.Delete_Point_Boxes(g["boxes"]) == int(g["ret_del"])
        xyz, ids = F.flatten()
        o = np.argsort(ids)
        assert np.array_equal(ids[o], g["final_ids"]) and np.array_equal(xyz[o], g["final_xyz"])
        gid, gsq, _, gcnt = F.Nearest_Search(g["queries"], 5, np.inf)
        assert np.array_equal(gcnt, g["q_cnt"]) and np.array_equal(gsq, g["q_sqd"]) and np.array_equal(gid, g["q_ids"])
    print("map ok")


def run_peer(pkg, n_map=60000, n_scan=4000):
    """Two ranks = two contexts of this process on cuda:0, one host thread each (see tests/test_multigpu_gpu.py)."""
    import flivo_scenes as scenes
    from flivo_parity import sharded_update_in_process
    world = scenes.make_world(seed=7, n_points=n_map, ds=0.5)
    scan = scenes.make_scan(world, seed=8, n_points=n_scan)
    sharded_update_in_process(pkg, world, scan, nranks=2, mode=pkg.MODE_V3, verbose=True)
    print("peer ok")


if __name__ == "__main__":
    pkg = ge.load_package()
    what = sys.argv[1:] or ["update", "map", "peer"]
    for w in what:
        {"update": run_update, "map": run_map, "peer": run_peer}[w](pkg)
