"""The C++ drop-in classes (fast-livo-noted_b200/host/: KD_TREE, LaserMapping) driven by a C++ caller written like
the reference's own call sites (tests/cpp/test_dropin.cpp).  CPU: it builds, links and refuses to run without a
device.  GPU: its results are compared with the oracle."""
import os
import struct
import subprocess

import numpy as np
import pytest

import flivo_parity as par
import flivo_scenes as scenes
from oracle import pyoracle as orc


def _binary(pkg):
    import importlib
    return importlib.import_module("fast_livo_noted_b200.build").build_host_test()


def test_dropin_builds_and_fails_loudly_without_gpu(pkg):
    import torch
    exe = _binary(pkg)
    r = subprocess.run([exe, "--no-gpu"], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0 and "constructed" in r.stdout
    else:
        assert r.returncode == 3 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_dropin_matches_oracle(pkg, tmp_path):
    world = scenes.make_world(seed=9, n_points=40000, ds=0.5)
    frames = [scenes.make_scan(world, seed=50 + i, n_points=3000) for i in range(3)]
    max_iter = 4
    path = os.path.join(tmp_path, "scene.bin")
    with open(path, "wb") as f:
        f.write(struct.pack("4i", len(world["points"]), 3000, len(frames), max_iter))
        f.write(struct.pack("d", world["ds"]))
        f.write(world["points"].astype(np.float32).tobytes())
        f.write(np.asarray(scenes.T_LI, np.float64).tobytes())
        for s in frames:
            f.write(np.asarray(s["x_prior"], np.float64).tobytes())
            f.write(s["body"].astype(np.float32).tobytes())
    r = subprocess.run([_binary(pkg), path], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = [l.split() for l in r.stdout.strip().splitlines()]
    kv = {}
    for l in lines:
        kv.setdefault(l[0], []).append(l[1:])
    pts = world["points"]
    assert int(kv["size"][0][0]) == len(pts) and kv["size"][0][2] == "1"
    nn = np.array(kv["nn5"][0], np.float64).reshape(5, 4)
    oid, osq, oxyz, ocnt, _ = par.oracle_knn(pts, pts[7:8], 5, np.inf)
    assert np.array_equal(nn[:, :3].astype(np.float32), oxyz[0]) and np.array_equal(nn[:, 3].astype(np.float32), osq[0])
    # frames: V1 update + map_incremental, against the oracle's V1 loop and Add_Points rule
    M = orc.OracleMap(world["ds"])
    M.build(pts)
    cur = pts.copy()
    for i, s in enumerate(frames):
        o = par.oracle_update(cur, s["body"], s["x_prior"], np.eye(18) * 1e-4, orc.MODE_V1, max_iter, v1_bound=pkg.v1_search_bound_sq(world["ds"]))
        fr = kv["frame"][i]
        got = dict(zip(fr[1::2], fr[2::2]))
        assert int(got["trips"]) == o["trips"] and int(got["searches"]) == o["searches"] and int(got["n_eff"]) == o["n_eff"]
        st = np.array(kv["state"][i], np.float64)
        assert np.abs(st[:9] - o["x"][:9]).max() < 1e-6 and np.abs(st[9:12] - o["x"][9:12]).max() < 1e-6
        wpts = orc.body_to_world(s["body"], orc.State.from_array(o["x"]))
        added = M.add_points(wpts, True)
        cur, _ = M.live()
        assert int(got["added"]) == added and int(got["size"]) == len(cur)
