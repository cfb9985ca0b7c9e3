"""Where the host-buffer (e2e) step spends its time: upload (repack + H2D + sort) vs update."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import __graft_entry__ as ge
import flivo_scenes as scenes
pkg = ge.load_package()
world = scenes.make_world_torch(3, 20_000_000, 0.2)
scan = scenes.make_scan_torch(world, 0, 260000, kind="ouster")
F = pkg.Flivo(downsample_size=0.2, initial_points=20_000_000, sort_scan=True)
F.Build(world["points"].cpu().numpy()); F.sync()
h = np.zeros((260000, 12), np.float32); h[:, :3] = scan["body"].cpu().numpy()
L = pkg.lib()
cfg = pkg.make_cfg(mode=0, max_iter=4)
P0 = np.eye(23) * 1e-4
for threads in (None,):
    up, upsync, upd = [], [], []
    for k in range(30):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        F._ck(L.flivo_scan_upload(F.h, h.ctypes.data, h.shape[0], 48)); F.n_scan = h.shape[0]
        t1 = time.perf_counter()
        F.sync()
        t2 = time.perf_counter()
        x, P, rep = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
        t3 = time.perf_counter()
        up.append(t1 - t0); upsync.append(t2 - t1); upd.append(t3 - t2)
    print("upload call %.1f us, drain after call %.1f us, update %.1f us (gpu %.1f us)" % (np.median(up) * 1e6, np.median(upsync) * 1e6, np.median(upd) * 1e6, rep.gpu_ms * 1e3))
    tot = []
    for k in range(30):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        F._ck(L.flivo_scan_upload(F.h, h.ctypes.data, h.shape[0], 48)); F.n_scan = h.shape[0]
        x, P, rep = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
        tot.append(time.perf_counter() - t0)
    print("back to back %.1f us" % (np.median(tot) * 1e6))
