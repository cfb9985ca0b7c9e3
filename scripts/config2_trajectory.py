#!/usr/bin/env python
"""BASELINE config 2 at size: dense Avia-like scans (100 k points) against a 5 M-point map, 100-scan trajectory, every scan:
iterated update, then map_incremental (flivo_map_add_scan = world transform with the posterior + Add_Points with
down-sampling; reference src/laser_mapping.cpp:165-251, 329-389, include/ikd-Tree/ikd_Tree.cpp:382-457).

Prints ONE JSON line: per-scan update / insert device+wall times, and — the parity statement of this configuration — the
live point SET of the GPU map after the whole trajectory against the reference's own ikd-Tree (oracle/_ref) that
received the same Add_Points calls (world points formed on the CPU from the GPU's posterior states: the transform is
bit-exact, tests/flivo_parity.py).  Usage: python scripts/config2_trajectory.py [--scans 100] [--no-cpu]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scans", type=int, default=100)
    ap.add_argument("--scan-points", type=int, default=100_000)
    ap.add_argument("--map-points", type=int, default=5_000_000)
    ap.add_argument("--ds", type=float, default=0.3)
    ap.add_argument("--mode", type=int, default=1, help="1 = V1 (the compiled path, what LaserMapping::Run runs), 0 = V3")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    json_fd = os.dup(1); os.dup2(2, 1); out = os.fdopen(json_fd, "w")   # the reference tree printf()s
    import torch
    import __graft_entry__ as ge
    import flivo_scenes as scenes
    pkg = ge.load_package()
    dev = "cuda:0"
    world = scenes.make_world_torch(seed=2, n_points=args.map_points, ds=args.ds, device=dev)
    wxyz = world["points"].cpu().numpy()
    # a drive along the streets: consecutive poses a few metres apart, looking along the direction of travel
    base = scenes.street_positions(world, 4)
    a, b = base[0], base[1] if len(base) > 1 else base[0] + np.array([20.0, 0, 0])
    poses = [a + (b - a) * (k / max(args.scans - 1, 1)) for k in range(args.scans)]
    npd = 18 if args.mode == 1 else 23
    P0 = np.eye(npd) * 1e-4
    cfg = pkg.make_cfg(mode=args.mode, max_iter=4, laser_point_cov=0.001)
    F = pkg.Flivo(device=0, downsample_size=args.ds, initial_points=args.map_points + args.scans * args.scan_points // 4, sort_scan=True)
    F.Build(wxyz)
    upd_ms, add_wall_ms, upd_wall_ms, added, trips, neff, posts, bodies = [], [], [], [], [], [], [], []
    for k in range(args.scans):
        s = scenes.make_scan_torch(world, seed=5000 + k, n_points=args.scan_points, kind="avia", pos=poses[k], device=dev)
        h = np.zeros((args.scan_points, 12), np.float32); h[:, :3] = s["body"].cpu().numpy()   # 48-byte PointXYZINormal records
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        F._ck(pkg.lib().flivo_scan_upload(F.h, h.ctypes.data, h.shape[0], 48)); F.n_scan = h.shape[0]
        x, P, rep = F.iekf_update(pkg.State.from_array(s["x_prior"]), P0, cfg)
        t1 = time.perf_counter()
        n_add = F.map_add_scan(x)
        F.sync()
        t2 = time.perf_counter()
        upd_ms.append(rep.gpu_ms); upd_wall_ms.append((t1 - t0) * 1e3); add_wall_ms.append((t2 - t1) * 1e3)
        added.append(int(n_add)); trips.append(int(rep.trips)); neff.append(int(rep.n_eff)); posts.append(x.as_array())
        if not args.no_cpu:
            bodies.append(np.ascontiguousarray(h[:, :3]))
    line = {"workload": "avia100k_map5M trajectory (BASELINE config 2)", "scans": args.scans, "scan_points": args.scan_points,
            "map_points_start": args.map_points, "map_points_end": int(F.size()), "downsample_size": args.ds, "mode": "V1" if args.mode else "V3",
            "update_device_ms": {"median": float(np.median(upd_ms)), "p99": float(np.percentile(upd_ms, 99)), "first": upd_ms[0]},
            "update_wall_ms_incl_upload": {"median": float(np.median(upd_wall_ms))},
            "map_incremental_wall_ms": {"median": float(np.median(add_wall_ms)), "p99": float(np.percentile(add_wall_ms, 99))},
            "points_added": {"total": int(np.sum(added)), "median": float(np.median(added))},
            "trips_per_scan": float(np.mean(trips)), "n_eff_median": float(np.median(neff)),
            "points_per_sec_update": float(args.scan_points * np.sum(trips) / (np.sum(upd_ms) * 1e-3))}
    if not args.no_cpu:
        from oracle import pyoracle as orc
        tree = orc.RefTree(args.ds)
        t0 = time.perf_counter(); tree.build(wxyz); tb = time.perf_counter() - t0
        cpu_add_ms, cpu_added = [], []
        for k in range(args.scans):
            w = orc.body_to_world(bodies[k], orc.State.from_array(posts[k]))
            t0 = time.perf_counter()
            cpu_added.append(int(tree.add_points(w, True)))
            cpu_add_ms.append((time.perf_counter() - t0) * 1e3)
        cx, cid = tree.flatten()
        gx, gid = F.flatten()
        og, oc = np.argsort(gid), np.argsort(cid)
        same = len(gid) == len(cid) and np.array_equal(gid[og], cid[oc]) and np.array_equal(gx[og], cx[oc])
        line["map_set_parity"] = {"against": "reference ikd_Tree.cpp (oracle/_ref) fed the same Add_Points calls", "identical": bool(same),
                                  "live_gpu": int(len(gid)), "live_cpu": int(len(cid)),
                                  "return_values_equal": bool(cpu_added == added),
                                  "ids_only_in_gpu": int(len(np.setdiff1d(gid, cid))), "ids_only_in_cpu": int(len(np.setdiff1d(cid, gid)))}
        line["cpu_add_points_ms"] = {"median": float(np.median(cpu_add_ms)), "tree_build_s": tb}
        tree.close()
    F.close()
    print(json.dumps(line), file=out, flush=True)


if __name__ == "__main__":
    main()
