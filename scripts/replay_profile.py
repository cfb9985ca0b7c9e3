"""Where a replay frame's update time goes (config 4: 24 k-point frames, map grown from the first scan): every tenth frame
runs in profile mode (per-trip events, plain launches) and prints the per-trip kernel times, the far-query statistics of
the first search trip and the per-CTA stamps; the other frames run the normal graph path and print their device time."""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=41)
    ap.add_argument("--pool", type=int, default=16)
    ap.add_argument("--scan-points", type=int, default=24000)
    ap.add_argument("--mode", type=int, default=1)
    a = ap.parse_args()
    torch.cuda.set_device(0)
    import __graft_entry__ as g
    pkg = g.load_package()
    import flivo_scenes as scenes
    w = scenes.make_world_torch(seed=1, n_points=1000000, ds=0.5, device="cuda")
    ends = scenes.street_positions(w, 2)
    poses = [ends[0] + (ends[1] - ends[0]) * (i / a.pool) for i in range(a.pool)]
    # feats_down_body is voxel-filtered (src/laser_mapping.cpp:129-130); leaf 0.3: the synthetic world shows only ~12 k distinct 0.5 m voxels from one pose
    pool = [{"body": s["body"].cpu().numpy(), "x_prior": np.asarray(s["x_prior"])}
            for s in scenes.make_replay_pool(w, poses, n_points=a.scan_points, leaf=0.3, kind="avia", device="cuda")]
    frames = [pool[k % a.pool] for k in range(a.frames)]
    npd = 18 if a.mode == 1 else 23
    P0 = np.eye(npd) * 1e-4
    cfg = pkg.make_cfg(mode=a.mode, max_iter=4)
    cfg_prof = pkg.make_cfg(mode=a.mode, max_iter=4, profile=True)
    with pkg.Flivo(device=0, downsample_size=0.5, initial_points=200000, sort_scan=True) as F:
        F.scan_upload(np.ascontiguousarray(frames[0]["body"], dtype=np.float32))
        wpts = F.h_share_model(pkg.State.from_array(frames[0]["x_prior"]), mode=a.mode, do_search=True, want=("world",))["world"]
        F.Build(wpts)
        for k in range(1, a.frames):
            prof = k % 10 == 0
            F.scan_upload(np.ascontiguousarray(frames[k]["body"], dtype=np.float32))
            x, _, rep = F.iekf_update(pkg.State.from_array(frames[k]["x_prior"]), P0, cfg_prof if prof else cfg)
            if prof:
                st = F.stats()
                trips = [(int(st.last_trip_searched[t]), round(float(st.last_fused_ms[t]) * 1e3, 1), round(float(st.last_solve_ms[t]) * 1e3, 1)) for t in range(rep.trips)]
                print(f"frame {k}: trips {rep.trips} searches {rep.searches} n_eff {rep.n_eff} map {F.size()} | per trip (searched, search+fused us, solve us): {trips} | far queries {int(st.last_far_queries)} candidates {int(st.last_candidates)}")
                ss = F.search_stamps().astype(np.int64)
                ss = ss[ss[:, 2] > 0]
                if len(ss):
                    t0 = ss[:, 0].min()
                    q = lambda v: "p50=%.1f p90=%.1f max=%.1f" % (np.percentile(v, 50), np.percentile(v, 90), v.max())
                    print(f"   search CTAs {len(ss)}: end of tile phase {q((ss[:, 1] - t0) / 1e3)} | end {q((ss[:, 2] - t0) / 1e3)} us")
                    ft = F.far_totals.astype(np.int64)
                    if ft[1]:
                        print(f"   far queries n={ft[1]} mean {ft[0] / ft[1] / 1e3:.1f} us max {ft[22] / 1e3:.1f} us; final radius histogram (cells) {dict((r, int(ft[3 + r])) for r in range(13) if ft[3 + r])}; mean us by radius {dict((r, round(ft[23 + r] / max(ft[3 + r], 1) / 1e3, 1)) for r in range(9) if ft[3 + r])}")
            else:
                print(f"frame {k}: device ms {getattr(rep, 'gpu_ms', 0.0):.3f} trips {rep.trips} searches {rep.searches}")
            F.map_add_scan(x)


if __name__ == "__main__":
    main()
