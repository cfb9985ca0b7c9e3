#!/usr/bin/env python
"""Generates the committed golden vectors under tests/golden/ by RUNNING THE REFERENCE'S OWN CODE in this
container: include/ikd-Tree/ikd_Tree.cpp compiled unmodified into oracle/_ref/libflivo_ref.so, and the header-only
include/ivox3d/ivox3d.h compiled unmodified into oracle/_ref/libivox_ref.so (oracle/Makefile).
The reference ships no tests or fixtures of its own, so these pin the kNN / Add_Points /
Delete_Point_Boxes half of the path and the iVox branch.  /root/reference is needed to (re)generate, never to run the tests.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import flivo_scenes as scenes  # noqa: E402
from oracle import pyoracle as orc  # noqa: E402


def knn_case():
    world = scenes.make_world(seed=21, n_points=20000, ds=0.5)
    pts = world["points"]
    rng = np.random.default_rng(5)
    q = pts[rng.choice(len(pts), 600, replace=False)] + rng.normal(0, 0.05, (600, 3)).astype(np.float32)
    q = np.concatenate([q, rng.uniform(-40, 40, (100, 3)).astype(np.float32)]).astype(np.float32)
    T = orc.RefTree(0.5)
    T.build(pts)
    out = dict(map=pts, queries=q)
    for tag, md in (("inf", np.inf), ("5", 5.0), ("03", 0.3)):
        ids, sqd, xyz, cnt = T.knn(q, 5, md, nthreads=1)
        out["ids_" + tag], out["sqd_" + tag], out["cnt_" + tag] = ids, sqd, cnt
    T.close()
    np.savez_compressed(os.path.join(HERE, "knn_ikdtree.npz"), **out)


def map_case():
    ds = 0.5
    world = scenes.make_world(seed=22, n_points=6000, ds=ds)
    base = world["points"]
    rng = np.random.default_rng(6)
    T = orc.RefTree(ds)
    T.build(base)
    out = dict(base=base, ds=np.float32(ds))
    # three Add_Points batches: re-observations of the same surfaces (downsample on), then one without downsampling
    for b in range(3):
        batch = (base[rng.choice(len(base), 2500)] + rng.normal(0, 0.08, (2500, 3))).astype(np.float32)
        fresh = rng.uniform(-30, 30, (500, 3)).astype(np.float32) * np.array([1, 1, 0.1], np.float32)
        batch = np.concatenate([batch, fresh]).astype(np.float32)
        out["batch%d" % b] = batch
        out["ret%d" % b] = np.int32(T.add_points(batch, True))
        xyz, ids = T.flatten()
        o = np.argsort(ids)
        out["live%d_xyz" % b], out["live%d_ids" % b] = xyz[o], ids[o]
    nods = rng.uniform(-5, 5, (300, 3)).astype(np.float32)
    out["batch_nods"] = nods
    out["ret_nods"] = np.int32(T.add_points(nods, False))
    boxes = np.array([[-10, -10, -1, 0, 0, 3], [5, 5, -1, 12.5, 9, 30]], np.float32)
    out["boxes"] = boxes
    out["ret_del"] = np.int32(T.delete_boxes(boxes))
    xyz, ids = T.flatten()
    o = np.argsort(ids)
    out["final_xyz"], out["final_ids"] = xyz[o], ids[o]
    # queries against the mutated tree
    q = (xyz[rng.choice(len(xyz), 400)] + rng.normal(0, 0.1, (400, 3))).astype(np.float32)
    out["queries"] = q
    kid, ksq, _, kcnt = T.knn(q, 5, np.inf, nthreads=1)
    out["q_ids"], out["q_sqd"], out["q_cnt"] = kid, ksq, kcnt
    T.close()
    np.savez_compressed(os.path.join(HERE, "map_ops_ikdtree.npz"), **out)


def lexsorted_rows(xyz, cnt):
    """Every row's first cnt points in lexicographic order, the rest zero (the reference returns nth_element's order)."""
    out = np.zeros_like(xyz)
    for i in range(len(cnt)):
        c = max(int(cnt[i]), 0)
        r = xyz[i, :c]
        out[i, :c] = r[np.lexsort((r[:, 2], r[:, 1], r[:, 0]))]
    return out


def ivox_case():
    """faster_lio::IVox<3, DEFAULT> itself: the reference's include/ivox3d headers compiled unmodified against the
    Eigen / pcl / glog stand-ins of oracle/shim (oracle/ivox_ref_wrap.cpp -> oracle/_ref/libivox_ref.so)."""
    res = 0.5
    world = scenes.make_world(seed=23, n_points=40000, ds=0.2)
    pts = world["points"]
    rng = np.random.default_rng(8)
    q = pts[rng.choice(len(pts), 700, replace=False)] + rng.normal(0, 0.04, (700, 3)).astype(np.float32)
    q = np.concatenate([q, rng.uniform(-30, 30, (100, 3)).astype(np.float32)]).astype(np.float32)
    out = dict(map=pts, queries=q, resolution=np.float32(res))
    for nearby in (1, 7, 19, 27):
        R = orc.RefIVox(res, nearby)
        for part in np.array_split(pts, 3):     # three AddPoints calls
            R.add_points(part)
        out["grids_%d" % nearby] = np.int64(R.num_grids())
        for tag, k, rng_m in (("k5", 5, 5.0), ("k3r02", 3, 0.2)):
            xyz, cnt = R.knn(q, k, rng_m)
            out["cnt_%d_%s" % (nearby, tag)] = cnt
            out["first_%d_%s" % (nearby, tag)] = xyz[:, 0].copy()
            out["set_%d_%s" % (nearby, tag)] = lexsorted_rows(xyz, cnt)
    # LRU eviction at a small capacity_ (ivox3d.h:262-265)
    cap = 300
    R = orc.RefIVox(res, 19, cap)
    batches = []
    for b in range(10):
        centre = rng.random(3) * 10 - 5
        batches.append((centre + rng.normal(0, 1.5, (600, 3))).astype(np.float32))
        R.add_points(batches[-1])
    ql = (rng.random((500, 3)) * 14 - 7).astype(np.float32)
    xyz, cnt = R.knn(ql, 5, 5.0)
    out.update(lru_capacity=np.int64(cap), lru_batches=np.stack(batches), lru_queries=ql, lru_grids=np.int64(R.num_grids()),
               lru_cnt=cnt, lru_set=lexsorted_rows(xyz, cnt))
    np.savez_compressed(os.path.join(HERE, "ivox_reference.npz"), **out)


if __name__ == "__main__":
    assert orc.have_ref(), "oracle/_ref/libflivo_ref.so missing: needs /root/reference"
    assert orc.have_ivox_ref(), "oracle/_ref/libivox_ref.so missing: needs /root/reference"
    knn_case()
    map_case()
    ivox_case()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
