#!/usr/bin/env python
"""bench.py — LIO measurement-update throughput of libflivo_b200.so on B200, next to the reference CPU path.

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]`; under torchrun one rank
per GPU.  Rank 0 prints ONE JSON line.

A "step" is one whole iterated measurement update (LaserMapping::Run step 5, src/laser_mapping.cpp:165-242 /
esekf::update_iterated_dyn_share_modified) of one synthetic scan against the resident map:
every h_share_model trip (world transform, exact 5-NN, esti_plane, gates, 1x12 rows, H^T H / H^T z) plus the
on-device solve, under the reference's own re-match policy.  metric = points/sec = N_scan * trips / time.

  value : scans already resident in HBM (flivo_scan_set_device + flivo_iekf_update), CUDA events on the ctx stream
  e2e   : the same through the reference-facing C-ABI calls with HOST buffers (flivo_scan_upload of the 48-byte
          PointXYZINormal AoS + flivo_iekf_update), wall clock, copies inside the timed region
  N > 1 : batched replay (BASELINE config 4 style): every rank holds a replica of the map and its own scans;
          no data-path collective; weak scaling.  `--workload mega` shards ONE scan over the ranks instead
          (config 5: allgather of the 96-double normal-equation block per trip).
The oracle (oracle/) is executed only by the cpu_baseline leg and by --impl reference.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

WORKLOADS = {
    # name: (scan points, map points, downsample size, scan kind, BASELINE.json config index)
    "avia24k_map1M": (24_000, 1_000_000, 0.5, "avia", 0),
    "avia100k_map5M": (100_000, 5_000_000, 0.3, "avia", 1),
    "ouster260k_map20M": (260_000, 20_000_000, 0.2, "ouster", 2),
    "mega4M_map100M": (4_000_000, 100_000_000, 0.2, "ouster", 4),
}
DEFAULT_WORKLOAD = "ouster260k_map20M"   # the configuration the >=50x target of north_star is quoted on
MAX_ITER = 4                             # "4 IEKF iterations" (BASELINE config 1); loop index -1 .. MAX_ITER-1
LASER_POINT_COV = 0.001
SEARCH_BYTES, REUSE_BYTES = 324, 120     # algorithmic bytes per (point, trip): SURVEY.md §8(d), DESIGN.md §roofline


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="flivo", choices=["flivo", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--scans", type=int, default=8, help="distinct scans (poses) rotated through the steps")
    ap.add_argument("--mode", type=int, default=0, help="0 = V3 (1x12 row, 23-DOF), 1 = V1 (compiled 18-DOF)")
    ap.add_argument("--force-search", type=int, default=0, help="1: kNN on every trip (worst case)")
    ap.add_argument("--sort-scan", type=int, default=1)
    ap.add_argument("--cell-ratio", type=int, default=0, help="search cell edge / downsample size (0 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-map-add", action="store_true", help="skip the informational map_incremental timing (it grows the map)")
    ap.add_argument("--no-ivox", action="store_true", help="skip the informational iVox-mode (SURVEY 8f rank 4) timing")
    ap.add_argument("--no-frontend", action="store_true", help="skip the informational front-end (undistort + voxel grid) timing")
    ap.add_argument("--cpu-steps", type=int, default=3)
    ap.add_argument("--shard-scan", action="store_true", help="N>1: shard ONE scan over the ranks (config 5) for the whole run")
    ap.add_argument("--no-sharded", action="store_true", help="N>1 replica run: skip the `sharded` block (config 5 on the same contexts)")
    ap.add_argument("--no-mega", action="store_true", help="N>1: skip the 4 M-point / 100 M-point part of the `sharded` block")
    ap.add_argument("--host-mem", choices=["registered", "pinned", "pageable"], default="registered",
                    help="where the e2e loop's host scans live: the node's own buffers registered with cudaHostRegister (default), "
                         "cudaHostAlloc memory (torch pin_memory), or plain pageable memory")
    ap.add_argument("--no-replay", action="store_true", help="skip the `replay` block (config 4 in miniature)")
    ap.add_argument("--replay-sequences", type=int, default=32, help="sequences PER RANK of the replay block")
    ap.add_argument("--replay-frames", type=int, default=150)
    ap.add_argument("--replay-concurrent", type=int, default=8)
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"], help="sharded mode: how the per-trip 96-double records travel")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "25"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                if t0 - 0.05 <= ts <= t1 + 0.15:
                    sm.append(float(f[0]))
                    for nm, v in zip(self.NAMES, f[3:7]):
                        if v.lower().startswith("active"):
                            reasons.add(nm)
                smax = float(f[1])
            except ValueError:
                continue
        out = {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "samples": len(sm), "reasons": sorted(reasons)}
        if not sm and self.rows:   # timed region shorter than the sampling period: report the closest sample and say so
            ts, line = min(self.rows, key=lambda r: abs(r[0] - t1))
            f = [x.strip() for x in line.split(",")]
            try:
                out["sm_mhz"] = float(f[0]); out["note"] = "no sample inside the %.2f s timed region; closest sample %.2f s away" % (t1 - t0, abs(ts - t1))
            except ValueError:
                pass
        return out


def make_workload(name, rank, n_scans, device):
    """Synthetic world + scans of BASELINE config `name` (tests/flivo_scenes.py; torch generators)."""
    import flivo_scenes as scenes
    n_scan, n_map, ds, kind, cfg_idx = WORKLOADS[name]
    world = scenes.make_world_torch(seed=cfg_idx + 1, n_points=n_map, ds=ds, device=device)
    poses = scenes.street_positions(world, n_scans * (rank + 1))[rank * n_scans:]
    scans = [scenes.make_scan_torch(world, seed=1000 * rank + i, n_points=n_scan, kind=kind, pos=poses[i], device=device)
             for i in range(n_scans)]
    return world, scans


# ---------------------------------------------------------------------------------------------------------
# reference CPU path: the reference's OWN ikd-Tree (oracle/_ref, compiled unmodified from /root/reference at
# build time) under the reference's OpenMP loop + the restated plane fit / Jacobian / solve (oracle/).
# ---------------------------------------------------------------------------------------------------------
def cpu_reference_run(world_xyz, scans_host, ds, mode, force_search, steps, warmup, threads, also_threads=None, trip_check=None):
    """-> dict(points_per_s, ms_per_step, trips, kind, cores, results[list of dict(x, n_eff, trips)]).
    also_threads: a second thread count timed on the same tree (MP_PROC_NUM = 4, what the reference's CMake
    picks on a host with more than 5 cores, CMakeLists.txt:33-36) -> keys alt_threads / alt_ms_per_step / alt_points_per_s."""
    from oracle import pyoracle as orc
    kind = "reference" if orc.have_ref() else "port"
    if kind != "reference":
        raise RuntimeError("oracle/_ref/libflivo_ref.so missing (built from /root/reference by __graft_entry__.build())")
    tree = orc.RefTree(ds)
    t0 = time.time()
    tree.build(world_xyz)
    build_s = time.time() - t0
    npd = 18 if mode == 1 else 23
    P0 = np.eye(npd) * 1e-4
    total_pts, total_s, results = 0, 0.0, []
    for it in range(warmup + steps):
        sc = scans_host[it % len(scans_host)]
        x = orc.State.from_array(sc["x_prior"])
        t0 = time.perf_counter()
        r = tree.iekf_update(sc["body"], x, P0, mode=mode, max_iter=MAX_ITER, laser_point_cov=LASER_POINT_COV,
                             nthreads=threads, force_search=bool(force_search))
        dt = time.perf_counter() - t0
        if it >= warmup:
            total_pts += len(sc["body"]) * r["trips"]
            total_s += dt
            results.append(dict(x=r["x"].as_array(), n_eff=r["n_eff"], trips=r["trips"], searches=r["searches"],
                                timings=[float(v) for v in r["timings"]], scan=it % len(scans_host)))
    alt = {}
    if also_threads and also_threads != threads:
        pts, secs = 0, 0.0
        for it in range(1 + max(1, min(steps, 2))):
            sc = scans_host[it % len(scans_host)]
            t0 = time.perf_counter()
            r = tree.iekf_update(sc["body"], orc.State.from_array(sc["x_prior"]), P0, mode=mode, max_iter=MAX_ITER,
                                 laser_point_cov=LASER_POINT_COV, nthreads=also_threads, force_search=bool(force_search))
            dt = time.perf_counter() - t0
            if it >= 1:
                pts += len(sc["body"]) * r["trips"]; secs += dt
        n_alt = max(1, min(steps, 2))
        alt = dict(alt_threads=also_threads, alt_ms_per_step=1e3 * secs / n_alt, alt_points_per_s=pts / secs)
    trip_parity = trip_check(tree) if trip_check else None   # (before Add_Points below changes the tree)
    # map_incremental of the last scan (src/laser_mapping.cpp:329-389): world transform with the posterior + Add_Points(true)
    sc = scans_host[results[-1]["scan"]]
    wpts = orc.body_to_world(sc["body"], orc.State.from_array(results[-1]["x"]))
    t0 = time.perf_counter()
    added = tree.add_points(wpts, True)
    add_ms = (time.perf_counter() - t0) * 1e3
    tree.close()
    return dict(points_per_s=total_pts / total_s, ms_per_step=1e3 * total_s / steps, kind=kind, cores=threads,
                build_s=build_s, results=results, add_points_ms=add_ms, add_points_added=int(added), trip_parity=trip_parity, **alt)


def bench_config(args):
    """The `config` of the JSON line: identical in both arms (the driver compares the dicts)."""
    n_scan, n_map, ds, kind, cfg_idx = WORKLOADS[args.workload]
    return {"workload": args.workload, "scan_points": n_scan, "map_points": n_map, "downsample_size": ds, "k": 5,
            "max_iter": MAX_ITER, "mode": "V1" if args.mode else "V3",
            "policy": "force_search" if args.force_search else "reference_rematch", "scans_rotated": args.scans}


def run_reference(args, rank, world_size):
    import torch
    if rank != 0:
        return None
    n_scan, n_map, ds, kind, cfg_idx = WORKLOADS[args.workload]
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    world, scans = make_workload(args.workload, 0, args.scans, dev)
    wxyz = world["points"].cpu().numpy()
    scans_host = [dict(body=s["body"].cpu().numpy(), x_prior=s["x_prior"]) for s in scans]
    del world, scans
    threads = os.cpu_count() or 1
    r = cpu_reference_run(wxyz, scans_host, ds, args.mode, args.force_search, args.steps, args.warmup, threads)
    sample = "%d full %d-pt scans x full %d-pt ikd-Tree, all trips" % (args.steps, n_scan, n_map)
    line = {
        "impl": "reference", "metric": "iekf_measurement_update_points_per_sec", "value": r["points_per_s"], "unit": "points/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 kNN/plane + f64 rows/solve",
        "data": "synthetic",
        "config": bench_config(args),
        "cpu_baseline": {"value": r["points_per_s"], "unit": "points/s", "cores": threads, "kind": r["kind"], "sample": sample,
                         "tree_build_s": r["build_s"]},
        "e2e": {"value": r["points_per_s"], "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    return line


def full_size_trip_parity(F, pkg, tree, body, x_prior, mode):
    """One searched h_share_model trip of the WHOLE bench scan on the GPU against the reference's own ikd-Tree
    (oracle/_ref Nearest_Search, k = 6 so that a float32 tie across rank 5 is seen) + the restated plane fit / gates:
    rows of neighbour ids that differ, point_selected_surf / effective-row flags that differ, ties, and the oracle's count
    of points whose plane test or s > 0.9 gate sits within rounding distance of its threshold (`borderline`)."""
    from oracle import pyoracle as orc
    import flivo_parity as par
    st = orc.State.from_array(x_prior)
    wpts = orc.body_to_world(body, st)
    oid, osq, oxyz, ocnt = tree.knn(wpts, 6, np.inf)
    ties = int(((ocnt >= 6) & (osq[:, 4] == osq[:, 5])).sum())
    cnt5 = np.minimum(ocnt, 5).astype(np.int32)
    d5 = np.where(cnt5 >= 5, osq[:, 4], np.inf).astype(np.float32)
    o = orc.h_share_model(body, st, np.ascontiguousarray(oxyz[:, :5]), cnt5, d5, True, mode, None)
    g = F.h_share_model(pkg.State.from_array(x_prior), mode, True, want=("selected", "eff", "nbr_id"))
    bound = pkg.v1_search_bound_sq(F.ds, F.cell_ratio) if mode == 1 else par.GATE_SQ      # V3 searches with the far-neighbour gate as its bound
    want = np.where((osq[:, :5] <= np.float32(bound)) & (np.arange(5)[None, :] < cnt5[:, None]), oid[:, :5], -1)
    bad_rows = np.any(want != g["nbr_id"], axis=1)
    # a row may legitimately differ only where a tie straddles rank 5 (ikd-Tree keeps the candidate it visited first)
    tie_rows = (ocnt >= 6) & (osq[:, 4] == osq[:, 5])
    return {"points": int(len(body)), "ids_mismatch_rows": int(bad_rows.sum()), "ids_mismatch_rows_not_explained_by_ties": int((bad_rows & ~tie_rows).sum()),
            "mask_mismatch": int((g["selected"] != o["mask"]).sum()), "eff_mismatch": int((g["eff"] != o["eff"]).sum()),
            "n_eff_gpu": int(g["n_eff"]), "n_eff_cpu": int(o["n_eff"]), "ties_across_rank5": ties, "borderline": int(o["borderline"]),
            "HTH_rel_err": float(par.rel_err_block(g["HTH"], o["HTH"])), "HTz_rel_err": float(par.rel_err_block(g["HTz"], o["HTz"])),
            "against": "reference ikd_Tree.cpp Nearest_Search (oracle/_ref) + restated esti_plane / gates / rows"}


def replay_block(args, pkg, rank, world_size, local_rank, dev, barrier):
    """BASELINE config 4 in miniature, inside the default run so that the scaling runs see it: independent sequences of
    24 k-point Avia-like frames, maps grown from the first scan, sequence s -> rank s mod G (sharding.replay_sequences:
    per frame flivo_scan_upload_async / commit, flivo_iekf_update_begin/_end, flivo_map_add_scan; one context and one host
    thread per sequence in flight, no collective).  Wall clock between two barriers, max over ranks."""
    import torch
    import torch.distributed as dist
    from importlib import import_module
    import flivo_scenes as scenes
    sh = import_module(pkg.__name__ + ".sharding")
    n_seq, n_frames, pool_n = args.replay_sequences * world_size, args.replay_frames, 16
    w = scenes.make_world_torch(seed=1, n_points=1_000_000, ds=0.5, device=dev)
    # a drive along a street: consecutive frames about 2 m apart (a sequence starts anywhere on the loop and wraps around)
    ends = scenes.street_positions(w, 2)
    poses = [ends[0] + (ends[1] - ends[0]) * (i / pool_n) for i in range(pool_n)]
    # feats_down_body is voxel-filtered (src/laser_mapping.cpp:129-130); leaf 0.3: the synthetic world shows only ~12 k distinct 0.5 m voxels from one pose
    pool = [{"body": s["body"].cpu().numpy(), "x_prior": np.asarray(s["x_prior"])}
            for s in scenes.make_replay_pool(w, poses, n_points=24_000, leaf=0.3, kind="avia", device=dev)]
    del w
    seqs = [[pool[(s + k) % pool_n] for k in range(n_frames)] for s in range(n_seq)]
    P18, cfg1 = np.eye(18) * 1e-4, pkg.make_cfg(mode=pkg.MODE_V1, max_iter=MAX_ITER, laser_point_cov=LASER_POINT_COV)
    sh.replay_sequences(pkg, [s[:6] for s in seqs[:world_size]], rank, world_size, P18, cfg1, ds=0.5, device=local_rank)   # warm-up
    barrier()
    t0 = time.perf_counter()
    # sequences in flight per rank.  (At N = 8 on one 32-core host the replay is host-bound either way: 8 in flight per rank
    # 10.6 k frames/s over the eight GPUs, 4 in flight 9.4 k, against 4.4 k on ONE GPU alone — every driving thread spins in
    # cudaStreamSynchronize while its frame runs.)
    conc = max(1, args.replay_concurrent)
    r = sh.replay_sequences(pkg, seqs, rank, world_size, P18, cfg1, ds=0.5, device=local_rank, concurrent=conc, reuse_contexts=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt, float(sum(v["work"] for v in r.values())), float(sum(v["updates"] for v in r.values())),
                      float(sum(v["device_ms"] for v in r.values()))], dtype=torch.float64, device=dev)
    if world_size > 1:
        mx, sm = t.clone(), t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dt, work, upd, busy = float(mx[0]), float(sm[1]), float(sm[2]), float(sm[3])
    else:
        dt, work, upd, busy = (float(v) for v in t)
    return {"workload": "config 4 in miniature: %d sequences x %d frames of 24 k voxel-filtered points (feats_down_body), V1, map grown from the first scan" % (n_seq, n_frames),
            "sequences": n_seq, "frames_per_sequence": n_frames, "concurrent_per_rank": conc,
            "points_per_sec": work / dt, "frames_per_sec": upd / dt, "ms_per_frame_per_rank": 1e3 * dt * world_size / max(upd, 1.0),
            "update_device_ms_mean": busy / max(upd, 1.0), "gpu_busy_updates": busy * 1e-3 / (dt * world_size),
            "note": "wall clock through the public API, host scans; every driving thread keeps one context for its sequences (Build replaces the map); no collective"}


def sharded_block(args, F, pkg, shared, P0, cfg, cfg_prof, rank, world_size, dev, barrier, wxyz_host, ds, workload=None, cpu_check=True):
    """BASELINE config 5 on the contexts of a replica run (N > 1): ONE scan split over the ranks (contiguous blocks of the
    caller's order, sharding.shard_range), every rank holding the whole map; per trip each rank's 96-double record is
    stored into every peer's exchange buffer from inside k_fused's solving CTA (NVLink peer memory), summed in rank
    order, solved redundantly.  Returns the `sharded` object of the JSON line (rank 0) — device time per update (max
    over ranks), exchange cost per trip from the solve's own globaltimer stamps, and the posterior checked against the
    reference-tree CPU path (oracle/_ref) on the WHOLE scan."""
    import torch
    import torch.distributed as dist
    from importlib import import_module
    shard_range = import_module(pkg.__name__ + ".sharding").shard_range
    handles = [None] * world_size
    dist.all_gather_object(handles, F.peer_export())
    F.peer_attach(handles, rank, world_size)
    parts = []
    for s in shared:
        lo, hi = shard_range(len(s["body"]), rank, world_size)
        b4 = torch.zeros((hi - lo, 4), dtype=torch.float32, device=dev)
        b4[:, :3] = s["body"][lo:hi]
        parts.append(b4)
    S = len(shared)

    def step(i, c=cfg):
        barrier()                                       # the ranks enter the update together (the exchange waits at most 2 s)
        F.scan_set_device(parts[i].data_ptr(), parts[i].shape[0])
        return F.iekf_update(pkg.State.from_array(shared[i]["x_prior"]), P0, c)

    for w in range(max(args.warmup, 3)):
        step(w % S)
    ms, trips = [], 0
    for k in range(args.steps):
        _, _, rp = step(k % S)
        ms.append(rp.gpu_ms); trips += rp.trips
    xphase, xwait = [], []
    for k in range(3):                                  # profile mode: plain launches, globaltimer stamps inside the solve
        _, _, rp = step(k % S, cfg_prof)
        st = F.solve_stamps(rp.trips).astype(np.int64)
        for t in range(rp.trips):
            if st[t][0] and st[t][1]:
                xphase.append((st[t][1] - st[t][0]) / 1e3)
            if st[t][10] and st[t][11]:
                xwait.append((st[t][11] - st[t][10]) / 1e3)
    tt = torch.tensor([sum(ms), float(np.mean(xphase)) if xphase else 0.0, float(np.mean(xwait)) if xwait else 0.0], dtype=torch.float64, device=dev)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    xs, Ps, reps = step(0)
    barrier()
    F.peer_detach()          # (the single-context cross-check below must not wait for peers)
    out = None
    if rank == 0:
        n_full = len(shared[0]["body"])
        out = {"workload": workload or args.workload, "scan_points": n_full, "ranks": world_size, "exchange": "peer memory (stores into the peers' HBM + release/acquire sequence flags inside k_fused)",
               "ms_per_step": float(tt[0]) / args.steps, "points_per_sec": n_full * trips / (float(tt[0]) * 1e-3), "trips_per_step": trips / args.steps,
               "exchange_us_per_trip": float(tt[1]), "exchange_wait_us_per_trip": float(tt[2]),
               "note": "ms_per_step: in-graph device time (staging + update), max over ranks; exchange_us_per_trip: record stores + flag + wait for all peers + rank-order sum (globaltimer stamps, profile mode, max over ranks); the wait share includes the ranks' arrival skew"}
        if not cpu_check:
            # too large for the CPU tree within the bench's time: the sharded posterior against a single-context update of
            # the WHOLE scan on this rank's own map replica (the single-GPU path is the one the oracle tests pin)
            full4 = torch.zeros((n_full, 4), dtype=torch.float32, device=dev)
            full4[:, :3] = shared[0]["body"]
            F.scan_set_device(full4.data_ptr(), n_full)
            x1, P1, rep1 = F.iekf_update(pkg.State.from_array(shared[0]["x_prior"]), P0, cfg)
            out["shard_check"] = {"against": "single-GPU update of the whole scan (same library, one context)",
                                  "x_max_abs_diff": float(np.abs(xs.as_array() - x1.as_array()).max()),
                                  "P_max_rel_diff": float(np.abs(np.asarray(Ps) - np.asarray(P1)).max() / np.abs(np.asarray(P1)).max()),
                                  "n_eff_sharded": int(reps.n_eff), "n_eff_single": int(rep1.n_eff),
                                  "trips_sharded": int(reps.trips), "trips_single": int(rep1.trips),
                                  "single_gpu_ms_per_step": float(rep1.gpu_ms)}
            del full4
        elif not args.no_cpu_baseline:
            from oracle import pyoracle as orc
            tree = orc.RefTree(ds)
            tree.build(wxyz_host)
            body = np.ascontiguousarray(shared[0]["body"].cpu().numpy())
            o = tree.iekf_update(body, orc.State.from_array(shared[0]["x_prior"]), P0, mode=args.mode, max_iter=MAX_ITER,
                                 laser_point_cov=LASER_POINT_COV, nthreads=os.cpu_count() or 1, force_search=bool(args.force_search))
            tree.close()
            out["shard_check"] = {"against": "reference ikd_Tree.cpp + restated loop on the whole scan (oracle/_ref), CPU",
                                  "x_max_abs_diff": float(np.abs(xs.as_array() - o["x"].as_array()).max()),
                                  "P_max_rel_diff": float(np.abs(np.asarray(Ps) - o["P"]).max() / np.abs(o["P"]).max()),
                                  "n_eff_sharded": int(reps.n_eff), "n_eff_cpu": int(o["n_eff"]),
                                  "trips_sharded": int(reps.trips), "trips_cpu": int(o["trips"])}
    barrier()
    return out


# ---------------------------------------------------------------------------------------------------------
# the product arm
# ---------------------------------------------------------------------------------------------------------
def run_flivo(args, rank, world_size, local_rank):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    pkg = ge.load_package()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libflivo_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = "cuda:%d" % local_rank
    n_scan, n_map, ds, kind, cfg_idx = WORKLOADS[args.workload]
    shard = bool(args.shard_scan) and world_size > 1
    world, scans = make_workload(args.workload, 0 if shard else rank, args.scans, dev)
    npd = 18 if args.mode == 1 else 23
    P0 = np.eye(npd) * 1e-4
    cfg = pkg.make_cfg(mode=args.mode, max_iter=MAX_ITER, laser_point_cov=LASER_POINT_COV, force_search=bool(args.force_search))
    cfg_prof = pkg.make_cfg(mode=args.mode, max_iter=MAX_ITER, laser_point_cov=LASER_POINT_COV, force_search=bool(args.force_search), profile=True)

    F = pkg.Flivo(device=local_rank, downsample_size=ds, cell_ratio=args.cell_ratio, initial_points=n_map, sort_scan=bool(args.sort_scan))
    # clocks of rank 0's GPU, sampled by nvidia-smi every 100 ms; started here so that it is up before the timed loops
    sampler = ClockSampler(local_rank if "CUDA_VISIBLE_DEVICES" not in os.environ else
                           os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local_rank]) if rank == 0 else None
    wxyz_host = world["points"].cpu().numpy()
    t0 = time.time()
    F.Build(wxyz_host)
    F.sync()
    map_build_s = time.time() - t0
    shared = None
    if world_size > 1 and not shard and not args.no_sharded:
        # the scans of the `sharded` block: the same two on every rank (rank 0's first poses), see sharded_block
        import flivo_scenes as scenes
        p0 = scenes.street_positions(world, args.scans)[:2]
        shared = [scenes.make_scan_torch(world, seed=i, n_points=n_scan, kind=kind, pos=p0[i], device=dev) for i in range(2)]
    del world
    if shard:   # config 5: contiguous block of every scan per rank + allgather of the normal-equation block per trip
        if args.exchange == "peer":   # records written straight into the peers' HBM by the solve kernel (NVLink)
            handles = [None] * world_size
            dist.all_gather_object(handles, F.peer_export())
            F.peer_attach(handles, rank, world_size)
        else:                         # NCCL all-gather between two solve-kernel launches
            uid = [pkg.Flivo.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(uid, src=0)
            F.comm_init(uid[0], rank, world_size)
        for s in scans:
            import importlib
            lo, hi = importlib.import_module("fast_livo_noted_b200.sharding").shard_range(len(s["body"]), rank, world_size)
            s["body_full"] = s["body"]
            s["body"] = s["body"][lo:hi].contiguous()
    # device-resident scans (float4) and the reference-layout host copies (48-byte PointXYZINormal AoS) for e2e
    dev_scans, host_scans, host_xyz, pinned_keep = [], [], [], []
    for s in scans:
        b4 = torch.zeros((len(s["body"]), 4), dtype=torch.float32, device=dev)
        b4[:, :3] = s["body"]
        dev_scans.append(b4)
        if args.host_mem == "pinned":
            hp = torch.zeros((len(s["body"]), 12), dtype=torch.float32).pin_memory()
            pinned_keep.append(hp)
            h = hp.numpy()
        else:
            h = np.zeros((len(s["body"]), 12), np.float32)
        h[:, :3] = s["body"].cpu().numpy()
        host_scans.append(h)
        host_xyz.append(np.ascontiguousarray(h[:, :3]))
    pts_per_scan = sum(len(s["body"]) for s in scans) / len(scans)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
    ext = torch.cuda.ExternalStream(F.stream(), device=dev)
    L = pkg.lib()
    import ctypes as C

    def barrier():
        torch.cuda.synchronize()
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident(i):
        F.scan_set_device(dev_scans[i].data_ptr(), dev_scans[i].shape[0])
        return F.iekf_update(pkg.State.from_array(scans[i]["x_prior"]), P0, cfg)

    def step_e2e(i):
        h = host_scans[i]
        F._ck(L.flivo_scan_upload(F.h, h.ctypes.data, h.shape[0], 48))
        F.n_scan = h.shape[0]
        return F.iekf_update(pkg.State.from_array(scans[i]["x_prior"]), P0, cfg)

    # ---- device-resident timing: per-step CUDA events on the ctx stream, L2 flushed (untimed) between steps ----
    S = len(scans)
    for w in range(args.warmup):
        step_resident(w % S)
    barrier()
    t_wall0 = time.time()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    trips_total, launches, fused_search_ms, fused_reuse_ms, solve_ms, last = 0, 0, [], [], [], None
    solve_us, graph_ms = [], []
    for k in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        ev[k][0].record(ext)
        x, P, rep = step_resident(k % S)
        ev[k][1].record(ext)
        trips_total += rep.trips
        launches += rep.kernel_launches   # own kernels only: k_scan_prepare, k_gather4, k_ctl_reset, (k_search, k_fused) per trip; the CUB radix-sort kernels of the staging graph are library code and not counted
        graph_ms.append(rep.gpu_ms)       # event-record NODES at the head of the staging graph and the tail of the update graph
        last = (k % S, x, P, rep)
    barrier()
    host_ev_ms = sorted(a.elapsed_time(b) for a, b in ev)   # events recorded from the host around the call (r1's method; moves when the host is busy)
    # a step = scan staging + update; its device time is read between two stamps the GPU takes itself (first node of the
    # staging graph, last node of the update graph), so host scheduling noise between the two graph launches cannot enter it
    step_ms = sorted(graph_ms)   # this rank's per-step device times (median + p99)
    dev_ms = sum(step_ms)
    tt = torch.tensor([dev_ms, float(trips_total) * pts_per_scan, float(launches)], dtype=torch.float64, device=dev)
    if world_size > 1:
        mx = tt.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, work, launches = float(mx[0]), float(sm[1]), int(sm[2])
    else:
        work = float(tt[1])
    value = work / (dev_ms * 1e-3)

    # ---- per-trip kernel durations for the roofline: same steps in profile mode (per-trip CUDA events, no graph) ----
    for k in range(min(args.steps, 8)):
        flush.zero_()
        torch.cuda.synchronize()
        F.scan_set_device(dev_scans[k % S].data_ptr(), dev_scans[k % S].shape[0])
        _, _, repp = F.iekf_update(pkg.State.from_array(scans[k % S]["x_prior"]), P0, cfg_prof)
        st = F.stats()
        for t in range(repp.trips):
            (fused_search_ms if st.last_trip_searched[t] else fused_reuse_ms).append(st.last_fused_ms[t])
            solve_ms.append(st.last_solve_ms[t])
        far_q, cand = int(st.last_far_queries), int(st.last_candidates)
        if k == min(args.steps, 8) - 1:
            ss = F.search_stamps().astype(np.int64)
            ss = ss[ss[:, 2] > 0]
            if len(ss):
                t0 = ss[:, 0].min()
                p1 = (ss[:, 1] - t0) / 1e3; en = (ss[:, 2] - t0) / 1e3; st0 = (ss[:, 0] - t0) / 1e3
                q = lambda a: "p50=%.1f p90=%.1f max=%.1f" % (np.percentile(a, 50), np.percentile(a, 90), a.max())
                print(f"[search CTAs, first search trip, us since first CTA start] ctas={len(ss)} start: {q(st0)} | end of tile phase: {q(p1)} | end: {q(en)} | far queries per CTA: max={int(ss[:, 3].max())} sum={int(ss[:, 3].sum())}", file=sys.stderr)
                ft = F.far_totals.astype(np.int64)
                if ft[1]:
                    print(f"[far queries, first search trip] n={ft[1]} mean {ft[0] / ft[1] / 1e3:.1f} us, {ft[2] / ft[1]:.2f} steps; final radius histogram (cells) {dict((r, int(ft[3 + r])) for r in range(13) if ft[3 + r])}; seeded candidates histogram {dict((c, int(ft[16 + c])) for c in range(6) if ft[16 + c])}; max {ft[22] / 1e3:.1f} us; mean us by final radius {dict((r, round(ft[23 + r] / max(ft[3 + r], 1) / 1e3, 1)) for r in range(9) if ft[3 + r])}", file=sys.stderr)
            stamps = F.solve_stamps(repp.trips)
            for t in range(repp.trips):
                row = stamps[t].astype(np.int64)
                d = [(int(row[i]) - int(row[i - 1])) / 1e3 if row[i] and row[i - 1] else None for i in range(1, 8)]
                tot = (int(row[7]) - int(row[0])) / 1e3 if row[7] and row[0] else None
                xw = (int(row[11]) - int(row[10])) / 1e3 if row[11] and row[10] else None
                print(f"[solve inside k_fused, trip {t}] us: exchange={d[0]} unpack={d[1]} aug={d[2]} gj={d[3]} gain={d[4]} plus+ctl={d[5]} cov|next-pre={d[6]} total={tot} peer-wait={xw}", file=sys.stderr)
                solve_us.append(tot)

    # ---- e2e: host buffers through the C ABI, wall clock, H2D + D2H inside the timed region ----
    # (a) serial (r1's number, kept for comparison): flivo_scan_upload of the 48-byte AoS from pageable memory, then
    #     flivo_iekf_update, one after the other; L2 flushed (untimed) between steps.
    for w in range(min(args.warmup, 2)):
        step_e2e(w % S)
    ser_s, ser_work = 0.0, 0.0
    for k in range(min(args.steps, 10)):
        flush.zero_()
        barrier()
        t0 = time.perf_counter()
        x2, P2, rep2 = step_e2e(k % S)
        ser_s += time.perf_counter() - t0
        ser_work += rep2.trips * host_scans[k % S].shape[0]
    e2e_serial = ser_work / ser_s
    # (b) pipelined (the headline e2e): the node hands one feats_down_body over per frame (src/laser_mapping.cpp:129-130,
    #     165-167).  While update k runs (flivo_iekf_update_begin), scan k+1 is staged (flivo_scan_upload_async: the
    #     48-byte records are repacked to float4 by host threads and travel on the copy stream — shipping the raw records
    #     was measured copy-bound, 12.5 MB per step over PCIe); flivo_iekf_update_end reads x, P and the report back.
    #     Wall clock over K back-to-back steps: K uploads (scan k+1 during step k), K updates, K result read-backs inside.
    #     No flush here: the rotated scans stand at different poses and together touch more of the map than L2 holds.
    states = [pkg.State.from_array(s["x_prior"]) for s in scans]

    # the four C-ABI calls of a frame, with their argument objects built once (what a C++ caller's loop looks like; the
    # convenience wrappers of the ctypes view copy the state and P on every call, which costs more than the calls)
    P_in = np.ascontiguousarray(P0, dtype=np.float64)
    P_out = np.zeros_like(P_in)
    x_out, rep_out = pkg.State(), pkg.Report()
    ptrs = [(C.c_void_p(h.ctypes.data), h.shape[0]) for h in host_scans]
    # the node's scan buffers registered once (cudaHostRegister through the ABI): the contract's "pinned host memory".  With
    # many host threads to itself the library still repacks the 48-byte records to float4 (a third of the PCIe bytes);
    # with few (8 ranks on one host) it sends the pinned records as they are.
    if args.host_mem == "registered":
        for h in host_scans:
            F._ck(L.flivo_host_register(C.c_void_p(h.ctypes.data), h.nbytes))
    hctx, p_in, p_out = F.h, C.c_void_p(P_in.ctypes.data), C.c_void_p(P_out.ctypes.data)
    ref_cfg, ref_x, ref_rep = C.byref(cfg), C.byref(x_out), C.byref(rep_out)
    ref_states = [C.byref(s) for s in states]

    def pipelined(n_steps):
        work = 0
        ck = F._ck
        ck(L.flivo_scan_upload_async(hctx, ptrs[0][0], ptrs[0][1], 48))
        barrier()
        t0 = time.perf_counter()
        for k in range(n_steps):
            i, j = k % S, (k + 1) % S
            ck(L.flivo_scan_commit(hctx))
            ck(L.flivo_iekf_update_begin(hctx, ref_states[i], p_in, ref_cfg))
            ck(L.flivo_scan_upload_async(hctx, ptrs[j][0], ptrs[j][1], 48))
            ck(L.flivo_iekf_update_end(hctx, ref_x, p_out, ref_rep))
            work += rep_out.trips * ptrs[i][1]
        dt = time.perf_counter() - t0
        ck(L.flivo_scan_commit(hctx))   # retire the scan staged during the last step
        return dt, float(work)

    pipelined(min(args.warmup, 3))
    e2e_s, e2e_work = pipelined(args.steps)
    # the clocks line covers the headline loops (device-resident, profile, e2e) and the sampler ends here: 40 NVML queries a second
    # take the driver's lock often enough to halve the launch-bound replay below (2.3 k frames/s with it, 6.3 k without)
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    # (replay first: measured at N = 2, the host side of the replay runs 7x slower in a process that has built and freed the
    # 20 M-node CPU reference tree AND staged the 100 M-point map before it; either alone does not matter)
    replay = replay_block(args, pkg, rank, world_size, local_rank, dev, barrier) if not args.no_replay else None
    sharded = sharded_block(args, F, pkg, shared, P0, cfg, cfg_prof, rank, world_size, dev, barrier, wxyz_host, ds) if shared else None
    if shared and not args.no_mega:
        # BASELINE config 5 proper: the 4 M-point mega-scan against the 100 M-point map (replicated: 5 GB of buckets per GPU)
        import flivo_scenes as scenes
        mn_scan, mn_map, mds, mkind, midx = WORKLOADS["mega4M_map100M"]
        mworld = scenes.make_world_torch(seed=midx + 1, n_points=mn_map, ds=mds, device=dev)
        mp0 = scenes.street_positions(mworld, 2)[:2]
        mshared = [scenes.make_scan_torch(mworld, seed=i, n_points=mn_scan, kind=mkind, pos=mp0[i], device=dev) for i in range(2)]
        Fm = pkg.Flivo(device=local_rank, downsample_size=mds, cell_ratio=args.cell_ratio, initial_points=mn_map, sort_scan=bool(args.sort_scan))
        Fm.Build(mworld["points"].cpu().numpy())
        del mworld
        mega = sharded_block(args, Fm, pkg, mshared, P0, cfg, cfg_prof, rank, world_size, dev, barrier, None, mds, workload="mega4M_map100M", cpu_check=False)
        Fm.close()
        del mshared
        if sharded is not None:
            sharded = {"scan260k": sharded, "mega4M_map100M": mega}
    te = torch.tensor([e2e_s, e2e_work], dtype=torch.float64, device=dev)
    if world_size > 1:
        mx = te.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = te.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        e2e_s, e2e_work = float(mx[0]), float(sm[1])
    # ---- front end, informational: raw scan (4 returns per down-sampled point, time-stamped) ->
    # undistort -> voxel grid on the device, timed with CUDA events; the same two steps in the CPU oracle beside it ----
    frontend = None
    if rank == 0 and world_size == 1 and not args.no_frontend:
        import flivo_frontend as fe
        g = torch.Generator(device=dev); g.manual_seed(1234)
        base = dev_scans[0][:, :3]
        raw = base.repeat_interleave(4, dim=0) + 0.02 * torch.randn((4 * base.shape[0], 3), device=dev, generator=g)
        n_raw = raw.shape[0]
        raw4 = torch.zeros((n_raw, 4), dtype=torch.float32, device=dev); raw4[:, :3] = raw
        t_ms = torch.linspace(0.0, 100.0, n_raw, device=dev, dtype=torch.float32)
        poses, rot_end, pos_end = fe.make_poses(9)
        R_LI, t_LI = np.eye(3), np.array([0.04165, 0.02326, -0.0284])
        leaf = max(ds, 0.05)
        und, dsm, n_out = [], [], 0
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        for k in range(6):
            F._ck(L.flivo_raw_set_device(F.h, C.c_void_p(raw4.data_ptr()), C.c_void_p(t_ms.data_ptr()), n_raw))
            torch.cuda.synchronize()
            with torch.cuda.stream(ext):
                e0.record(ext)
                F.raw_undistort(poses, rot_end, pos_end, R_LI, t_LI)
                e1.record(ext)
                n_out = F.raw_downsample(leaf)
                e2.record(ext)
            torch.cuda.synchronize()
            if k >= 2:
                und.append(e0.elapsed_time(e1)); dsm.append(e1.elapsed_time(e2))
        frontend = {"raw_points": int(n_raw), "leaf": leaf, "scan_points_out": int(n_out), "undistort_ms": float(np.median(und)),
                    "voxel_grid_ms": float(np.median(dsm)), "note": "voxel_grid_ms: PCL-equivalent voxel grid on the device; the result is the current scan (its Morton staging runs with the next update)"}
        if not args.no_cpu_baseline:
            from oracle import pyoracle as orc
            rh = raw.cpu().numpy(); th = t_ms.cpu().numpy()
            t0 = time.perf_counter(); uh = orc.undistort(rh, th, poses, rot_end, pos_end, R_LI, t_LI); t1 = time.perf_counter()
            orc.voxel_grid(np.c_[uh, np.zeros(len(uh), np.float32)].astype(np.float32), leaf, t_ms=th); t2 = time.perf_counter()
            frontend["cpu_undistort_ms"] = (t1 - t0) * 1e3
            frontend["cpu_voxel_grid_ms"] = (t2 - t1) * 1e3
            frontend["cpu_kind"] = "port (oracle restatement, 1 thread)"
        del raw, raw4, t_ms
    # ---- map_incremental (flivo_map_add_scan: world transform with the posterior + Add_Points with down-sampling),
    # informational, BASELINE config 2 style: each scan is inserted once, after its update ----
    map_add = None
    if rank == 0 and world_size == 1 and not args.no_map_add:
        ms, added = [], []
        for i in range(S):
            F.scan_set_device(dev_scans[i].data_ptr(), dev_scans[i].shape[0])
            xi, _, _ = F.iekf_update(pkg.State.from_array(scans[i]["x_prior"]), P0, cfg)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            added.append(F.map_add_scan(xi))
            F.sync()
            ms.append((time.perf_counter() - t0) * 1e3)
        map_add = {"ms_per_scan_median": float(np.median(ms)), "ms_per_scan": [float(v) for v in ms], "points_added": [int(a) for a in added],
                   "scan_points": int(pts_per_scan), "note": "wall clock around flivo_map_add_scan + sync; the map grows by the added points"}
    # ---- iVox-compatible mode (the local map the reference compiles by default), informational:
    # the same scans, V1 (compiled 18-DOF) update, against an iVox map (resolution 0.5, NEARBY18) holding the map points
    # within lidar range of the sensor; the CPU arm is the restated IVox + loop on all host threads (the restatement is pinned on the reference's compiled header: tests/test_ivox_cpu.py).
    ivox = None
    if rank == 0 and world_size == 1 and not args.no_ivox:
        import flivo_parity as par
        from oracle import pyoracle as orc
        near = wxyz_host[(np.abs(wxyz_host[:, 0]) < 130.0) & (np.abs(wxyz_host[:, 1]) < 130.0)]
        res, nearby = 0.5, 19
        Fv = pkg.Flivo(device=local_rank, downsample_size=res, initial_points=len(near), sort_scan=bool(args.sort_scan), ivox_nearby=nearby)
        Fv.AddPoints(near)
        Fv.sync()
        P18 = np.eye(18) * 1e-4
        cfg1 = pkg.make_cfg(mode=pkg.MODE_V1, max_iter=MAX_ITER, laser_point_cov=LASER_POINT_COV)
        extv = torch.cuda.ExternalStream(Fv.stream(), device=dev)
        ms, trips_v, rep_v, x_v = [], 0, None, None
        for k in range(args.warmup + min(args.steps, 10)):
            i = k % S
            flush.zero_()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(extv)
            Fv.scan_set_device(dev_scans[i].data_ptr(), dev_scans[i].shape[0])
            x_v, _, rep_v = Fv.iekf_update(pkg.State.from_array(scans[i]["x_prior"]), P18, cfg1)
            e1.record(extv)
            torch.cuda.synchronize()
            if k >= args.warmup:
                ms.append(e0.elapsed_time(e1)); trips_v += rep_v.trips
        i_last = (args.warmup + min(args.steps, 10) - 1) % S
        t0 = time.perf_counter()
        n_add = Fv.map_add_scan(x_v)
        Fv.sync()
        add_ms = (time.perf_counter() - t0) * 1e3
        ivox = {"resolution": res, "nearby": "NEARBY18", "map_points": int(len(near)), "mode": "V1", "ms_per_update_median": float(np.median(ms)),
                "points_per_sec": float(pts_per_scan * trips_v / (sum(ms) * 1e-3)), "trips_per_update": trips_v / len(ms),
                "map_incremental_ms": add_ms, "map_incremental_added": int(n_add)}
        if not args.no_cpu_baseline:
            v = orc.OracleIVox(res, nearby)
            t0 = time.perf_counter(); v.add_points(near); t_build = time.perf_counter() - t0
            body = host_xyz[i_last]
            t0 = time.perf_counter()
            o = par.oracle_update(None, body, scans[i_last]["x_prior"], P18, orc.MODE_V1, MAX_ITER, LASER_POINT_COV, None, False, ivox=v)
            t_cpu = time.perf_counter() - t0
            ivox.update(cpu_ms_per_update=t_cpu * 1e3, cpu_map_build_s=t_build, cpu_kind="port (restated IVox + loop, OpenMP search, %d threads)" % (os.cpu_count() or 1),
                        parity={"x_max_abs_diff": float(np.abs(x_v.as_array() - o["x"]).max()), "n_eff_gpu": int(rep_v.n_eff), "n_eff_cpu": int(o["n_eff"]),
                                "trips_gpu": int(rep_v.trips), "trips_cpu": int(o["trips"]), "distance_ties": int(o["ties"])})
            del v
        del Fv
    # sharded mode: the posterior of the sharded update must be the single-GPU posterior of the whole scan
    shard_check = None
    if shard:
        barrier()
        xs, Ps, reps = step_resident(0)
        barrier()
        if rank == 0:
            F1 = pkg.Flivo(device=local_rank, downsample_size=ds, cell_ratio=args.cell_ratio, initial_points=n_map, sort_scan=bool(args.sort_scan))
            F1.Build(wxyz_host)
            full = np.ascontiguousarray(scans[0]["body_full"].cpu().numpy())
            F1.scan_upload(full)
            x1, P1, rep1 = F1.iekf_update(pkg.State.from_array(scans[0]["x_prior"]), P0, cfg)
            shard_check = {"x_max_abs_diff_vs_single_gpu": float(np.abs(xs.as_array() - x1.as_array()).max()),
                           "P_max_abs_diff_vs_single_gpu": float(np.abs(np.asarray(Ps) - np.asarray(P1)).max()),
                           "n_eff_sharded": int(reps.n_eff), "n_eff_single": int(rep1.n_eff),
                           "trips_sharded": int(reps.trips), "trips_single": int(rep1.trips)}
            del F1
        barrier()
    # xyz as float4 (repacked from the 48-byte records on the host) — or the pinned 48-byte records as they are when the process
    # has few host threads to itself (the library's rule, flivo_scan_upload_async) — + the job / state / P block
    raw_env = os.environ.get("FLIVO_RAW_DMA")
    nt_host = int(os.environ.get("FLIVO_HOST_THREADS") or min(16, max(1, len(os.sched_getaffinity(0)) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1"))))))
    few = raw_env is None and nt_host <= 8 and pts_per_scan >= 65536
    raw_dma = args.host_mem != "pageable" and (raw_env in ("1", "2") or few)
    split_pct = 0 if raw_env == "1" else min(100, max(0, int(os.environ.get("FLIVO_RAW_SPLIT", "50" if nt_host <= 4 else "80"))))   # per cent of a scan repacked by the host
    bytes_per_point = (16 * split_pct + 48 * (100 - split_pct)) / 100.0 if raw_dma else 16
    h2d = int(pts_per_scan * bytes_per_point + 36 * 8 + npd * npd * 8 + 512)
    d2h = int(36 * 8 + npd * npd * 8 + 512)
    return dict(F=F, pkg=pkg, value=value, dev_ms=dev_ms, e2e=e2e_work / e2e_s, h2d=h2d, d2h=d2h, raw_dma=raw_dma, split_pct=split_pct, nt_host=nt_host, host_mem=args.host_mem, launches=launches, clocks=clocks,
                fused_search_ms=fused_search_ms, fused_reuse_ms=fused_reuse_ms, solve_ms=solve_ms, last=last, scans=scans,
                host_xyz=host_xyz, wxyz_host=wxyz_host, map_build_s=map_build_s, pts_per_scan=pts_per_scan,
                trips_per_step=trips_total / args.steps, shard=shard, shard_check=shard_check, frontend=frontend, map_add=map_add, ivox=ivox,
                step_ms_median=float(np.median(step_ms)), step_ms_p99=float(np.percentile(step_ms, 99)), step_ms_min=float(step_ms[0]),
                host_ev_ms_mean=float(np.mean(host_ev_ms)), sharded=sharded, replay=replay, solve_us=[v for v in solve_us if v is not None], e2e_serial=e2e_serial,
                search_stats={"far_queries_last_step": far_q, "candidates_last_step": cand})


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def main():
    args = parse_args()
    # the reference's ikd-Tree printf()s to stdout; keep fd 1 clean for the ONE JSON line
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    out = os.fdopen(json_fd, "w")
    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        line = run_reference(args, rank, world_size)
        if line is not None:
            print(json.dumps(line), file=out, flush=True)
        return 0
    import torch
    import torch.distributed as dist
    one_device = os.environ.get("FLIVO_BENCH_ONE_DEVICE") == "1"   # tests: every rank on cuda:0 (cudaIpc across processes on one GPU)
    if one_device:
        local_rank = 0
    if world_size > 1:
        torch.cuda.set_device(local_rank)
        if one_device:
            dist.init_process_group("gloo")     # NCCL refuses two ranks on one device; only rendezvous / timings travel here
        else:
            dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % local_rank))
    r = run_flivo(args, rank, world_size, local_rank)
    n_scan, n_map, ds, kind, cfg_idx = WORKLOADS[args.workload]
    peak, peak_src = peaks()
    # roofline of the dominant kernel (k_fused on a search trip): algorithmic bytes / measured launch duration
    # a trip = k_search + k_fused, and k_fused's last CTA also runs the trip's solve (a serial ~18 us tail that touches no
    # scan data); the roofline is quoted for the data-parallel pair, so the solve's own stamped duration is taken off
    solve_ms_in = (float(np.mean(r["solve_us"])) * 1e-3) if r["solve_us"] else 0.0
    ms_s = (float(np.mean(r["fused_search_ms"])) - solve_ms_in) if r["fused_search_ms"] else None
    ms_r = (float(np.mean(r["fused_reuse_ms"])) - solve_ms_in) if r["fused_reuse_ms"] else None
    n_local = r["pts_per_scan"]
    ach = SEARCH_BYTES * n_local / (ms_s * 1e-3) / 1e9 if ms_s else None
    roof = {"bound": "hbm", "kernel": "k_search + k_fused of a search trip, without the trip's solve (which runs inside k_fused's last CTA; its stamped duration is subtracted)", "achieved": ach, "peak": peak, "unit": "GB/s",
            "frac": (ach / peak) if ach else None, "traffic": None, "peak_source": peak_src,
            "bytes_per_point_trip": SEARCH_BYTES, "points_per_launch": n_local, "avg_launch_ms": ms_s,
            "reuse_trip": {"bytes_per_point_trip": REUSE_BYTES, "avg_launch_ms": ms_r,
                           "achieved": (REUSE_BYTES * n_local / (ms_r * 1e-3) / 1e9) if ms_r else None},
            "k_fused_incl_solve_avg_ms": float(np.mean(r["solve_ms"])) if r["solve_ms"] else None,
            "solve_inside_k_fused_us": (float(np.mean(r["solve_us"])) if r["solve_us"] else None)}
    tp = os.path.join(ROOT, "profiles", "traffic.json")   # dram bytes per launch from the committed ncu --set full capture
    if os.path.exists(tp):
        try:
            roof["traffic"] = json.load(open(tp)).get(args.workload)
        except Exception:
            pass
    cpu = None
    parity = None
    if rank == 0 and world_size == 1 and not args.no_cpu_baseline:
        scans_host = [dict(body=r["host_xyz"][i], x_prior=r["scans"][i]["x_prior"]) for i in range(len(r["scans"]))]
        threads = os.cpu_count() or 1
        F_, pkg_ = r["F"], r["pkg"]

        def trip_check(tree):   # whole scan 0, one searched trip: ids / masks of all points against the reference tree
            F_.Build(r["wxyz_host"])        # (the informational map_incremental block above has grown the GPU map)
            F_.scan_upload(r["host_xyz"][0])
            return full_size_trip_parity(F_, pkg_, tree, r["host_xyz"][0], r["scans"][0]["x_prior"], args.mode)

        c = cpu_reference_run(r["wxyz_host"], scans_host, ds, args.mode, args.force_search, args.cpu_steps, 1, threads, also_threads=4, trip_check=trip_check)
        cpu = {"value": c["points_per_s"], "unit": "points/s", "cores": threads, "kind": c["kind"],
               "mp_proc_num_4": ({"value": c["alt_points_per_s"], "ms_per_scan": c["alt_ms_per_step"], "cores": c["alt_threads"],
                                  "note": "MP_PROC_NUM = 4: the thread count the reference's own CMake picks (CMakeLists.txt:33-36); "
                                          "the headline uses the faster all-cores figure"} if "alt_threads" in c else None),
               "sample": "%d full %d-pt scans x full %d-pt ikd-Tree (tree build %.1f s untimed), all trips" % (args.cpu_steps, n_scan, n_map, c["build_s"]),
               "ms_per_scan": c["ms_per_step"], "stage_s_last": c["results"][-1]["timings"],
               "add_points_ms": c["add_points_ms"], "add_points_added": c["add_points_added"]}
        # parity side-check on the bench workload itself: GPU posterior of scan j vs the reference CPU path's
        j, xg, Pg, rep = r["last"]
        cands = [q for q in c["results"] if q["scan"] == j]
        if not cands:
            F, pkg = r["F"], r["pkg"]
            j = c["results"][-1]["scan"]
            F.scan_upload(r["host_xyz"][j])
            cfg = pkg.make_cfg(mode=args.mode, max_iter=MAX_ITER, laser_point_cov=LASER_POINT_COV, force_search=bool(args.force_search))
            xg, Pg, rep = F.iekf_update(pkg.State.from_array(r["scans"][j]["x_prior"]), np.eye(18 if args.mode == 1 else 23) * 1e-4, cfg)
            cands = [c["results"][-1]]
        q = cands[-1]
        parity = {"scan": j, "x_max_abs_diff": float(np.abs(xg.as_array() - q["x"]).max()), "n_eff_gpu": int(rep.n_eff),
                  "n_eff_cpu": int(q["n_eff"]), "trips_gpu": int(rep.trips), "trips_cpu": int(q["trips"]),
                  "first_trip_all_points": c["trip_parity"]}
    if rank == 0:
        line = {
            "metric": "iekf_measurement_update_points_per_sec", "value": r["value"], "unit": "points/s", "n_gpus": world_size,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["dev_ms"] / args.steps, "higher_is_better": True,
            "scaling": "strong" if r["shard"] else "weak", "vs_baseline": None, "dtype": "f32 kNN/plane + f64 rows/solve", "data": "synthetic",
            "config": bench_config(args),
            "run": {"trips_per_step": r["trips_per_step"], "sort_scan": bool(args.sort_scan),
                    "l2": "value: 512 MiB buffer written between timed steps (untimed). e2e: no flush, %d rotated scans at different poses (map %d pts = %.1f GB of buckets > 126 MB L2)" % (args.scans, n_map, n_map * 40 / 1e9),
                    "timing": "value: per step, CUDA event-record nodes at the head of the staging graph and the tail of the update graph, summed over the steps, max over ranks. e2e: wall clock over the K pipelined steps, max over ranks",
                    "multi_gpu": (("one scan sharded over ranks; per trip 96 doubles per rank, " + ("stored into the peers' HBM from inside k_fused's solving CTA (NVLink peer memory)" if args.exchange == "peer" else "ncclAllGather")) if r["shard"] else
                                  "map replica + own scans per rank, no collective")},
            "latency_ms_per_scan": r["dev_ms"] / args.steps,
            "latency_ms_per_scan_stats": {"median": r["step_ms_median"], "p99": r["step_ms_p99"], "min": r["step_ms_min"],
                                          "host_recorded_events_mean": r["host_ev_ms_mean"],
                                          "note": "per-step device times of rank 0 over the timed steps (in-graph event nodes); host_recorded_events_mean = the same steps between events recorded from the host around the call"},
            "e2e": {"value": r["e2e"], "unit": "points/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                    "how": "flivo_scan_upload_async of the next 48-byte-AoS scan (" + r["host_mem"] + " host buffer; " + (("%d %% of the scan repacked to float4 by the host, the rest of the pinned records sent as they are, both on the copy stream" % r["split_pct"]) if r["raw_dma"] else "host repack to float4 + copy stream") + ") during flivo_iekf_update_begin/_end of the current one",
                    "host_threads_per_rank": r["nt_host"],
                    "serial_value": r["e2e_serial"], "serial_how": "flivo_scan_upload (pageable buffer, host repack) then flivo_iekf_update, L2 flushed between steps"},
            "gpu_launches": r["launches"], "clocks": r["clocks"], "roofline": roof, "cpu_baseline": cpu, "parity_check": parity, "shard_check": r["shard_check"], "sharded": r["sharded"], "replay": r["replay"], "frontend": r["frontend"], "map_add": r["map_add"], "ivox": r["ivox"],
            "map_build_s": r["map_build_s"], "search_stats": r["search_stats"],
        }
        print(json.dumps(line), file=out, flush=True)
    r["F"].close()
    if world_size > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
