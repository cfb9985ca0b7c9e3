"""BASELINE config 4 in miniature: S independent sequences of F Avia-like frames (24 k points after filter_size_surf
= 0.5, map growing from the first scan), sequence s -> rank s mod G, no collective (sharding.replay_sequences).
Run plain (1 GPU) or under torchrun (one rank per GPU).  Prints ONE JSON line on rank 0: whole-job points/s =
sum over ranks of (trips x scan points) / max over ranks of the wall time between two barriers.
The frames come from a pool of P scans along a street of the config-1 world; sequence s starts at pool[s mod P]."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sequences", type=int, default=64)
    ap.add_argument("--frames", type=int, default=20)
    ap.add_argument("--pool", type=int, default=16)
    ap.add_argument("--scan-points", type=int, default=24000)
    ap.add_argument("--mode", type=int, default=1)
    ap.add_argument("--concurrent", type=int, default=1, help="sequences of a rank driven at a time (host threads)")
    ap.add_argument("--reuse", type=int, default=1, help="1: every driving thread keeps one context for all its sequences (Build replaces the map); 0: one context per sequence")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
    import importlib
    import __graft_entry__ as g
    pkg = g.load_package()
    sh = importlib.import_module("fast_livo_noted_b200.sharding")
    import flivo_scenes as scenes
    w = scenes.make_world_torch(seed=1, n_points=1000000, ds=0.5, device="cuda")
    ends = scenes.street_positions(w, 2)      # a drive along a street: consecutive frames (30 m / pool) apart
    poses = [ends[0] + (ends[1] - ends[0]) * (i / a.pool) for i in range(a.pool)]
    # feats_down_body is voxel-filtered (src/laser_mapping.cpp:129-130); leaf 0.3: the synthetic world shows only ~12 k distinct 0.5 m voxels from one pose
    pool = [{"body": s["body"].cpu().numpy(), "x_prior": np.asarray(s["x_prior"])}
            for s in scenes.make_replay_pool(w, poses, n_points=a.scan_points, leaf=0.3, kind="avia", device="cuda")]
    seqs = [[pool[(s + k) % a.pool] for k in range(a.frames)] for s in range(a.sequences)]
    npd = 18 if a.mode == 1 else 23
    P0, cfg = np.eye(npd) * 1e-4, pkg.make_cfg(mode=a.mode, max_iter=4)
    sh.replay_sequences(pkg, seqs[:world], rank, world, P0, cfg, ds=0.5, device=dev)      # warm-up: one sequence per rank
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    r = sh.replay_sequences(pkg, seqs, rank, world, P0, cfg, ds=0.5, device=dev, concurrent=a.concurrent, reuse_contexts=bool(a.reuse))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt, float(sum(v["work"] for v in r.values())), float(sum(v["updates"] for v in r.values()))], dtype=torch.float64, device=torch.device("cuda", dev))
    if world > 1:
        mx, sm = t.clone(), t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dt, work, upd = float(mx[0]), float(sm[1]), float(sm[2])
    else:
        dt, work, upd = float(t[0]), float(t[1]), float(t[2])
    if rank == 0:
        print(json.dumps({"metric": "replay_points_per_sec", "value": work / dt, "unit": "points/s", "n_gpus": world,
                          "sequences": a.sequences, "frames_per_sequence": a.frames, "scan_points": a.scan_points, "concurrent_per_rank": a.concurrent, "updates": int(upd),
                          "ms_per_frame_wall": 1e3 * dt * world / max(upd, 1.0), "mode": "V1" if a.mode == 1 else "V3",
                          "frames_per_sec": upd / dt, "gpu_busy_updates_rank0": sum(v["device_ms"] for v in r.values()) * 1e-3 / dt,
                          "map_points_seq0": r[min(r)]["map_points"], "scaling": "weak per sequence, no collective",
                          "note": "wall clock through the public API (host scans, upload, update, map_incremental, per-sequence context create/destroy)"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
