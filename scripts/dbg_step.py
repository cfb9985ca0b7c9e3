"""Device-side split of one resident step: scan staging (keys + sort + gather) vs the update, CUDA events on the ctx stream."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import __graft_entry__ as ge
import flivo_scenes as scenes
pkg = ge.load_package()
world = scenes.make_world_torch(3, 20_000_000, 0.2)
scans = [scenes.make_scan_torch(world, i, 260000, kind="ouster") for i in range(2)]
F = pkg.Flivo(downsample_size=0.2, initial_points=20_000_000, sort_scan=True)
F.Build(world["points"].cpu().numpy()); F.sync()
dev = []
for s in scans:
    b4 = torch.zeros((260000, 4), dtype=torch.float32, device="cuda"); b4[:, :3] = s["body"]; dev.append(b4)
ext = torch.cuda.ExternalStream(F.stream(), device="cuda")
cfg = pkg.make_cfg(mode=0, max_iter=4)
P0 = np.eye(23) * 1e-4
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
a, b, host_a, host_b = [], [], [], []
for k in range(24):
    i = k % 2
    flush.zero_(); torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(ext)
    t0 = time.perf_counter()
    F.scan_set_device(dev[i].data_ptr(), 260000)
    t1 = time.perf_counter()
    e1.record(ext)
    x, P, rep = F.iekf_update(pkg.State.from_array(scans[i]["x_prior"]), P0, cfg)
    t2 = time.perf_counter()
    e2.record(ext)
    torch.cuda.synchronize()
    if k >= 4:
        a.append(e0.elapsed_time(e1) * 1e3); b.append(e1.elapsed_time(e2) * 1e3); host_a.append((t1 - t0) * 1e6); host_b.append((t2 - t1) * 1e6)
print("staging span %.1f us (host call %.1f us) | update span %.1f us (host call %.1f us, graph-internal gpu_ms %.1f us)" %
      (np.median(a), np.median(host_a), np.median(b), np.median(host_b), rep.gpu_ms * 1e3))
