import sys, os, ctypes as C
sys.path.insert(0, os.getcwd()); sys.path.insert(0, "tests")
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
import flivo_scenes as scenes
world = scenes.make_world(seed=7, n_points=60000, ds=0.5)
scan = scenes.make_scan(world, seed=8, n_points=4000)
F = pkg.Flivo(downsample_size=0.5, initial_points=60000, sort_scan=True)
F.Build(world["points"]); F.scan_upload(scan["body"])
L = pkg.lib()
L.flivo_debug_sum.argtypes = [C.c_void_p, C.c_void_p]
for it in range(3):
    x, P, rep = F.iekf_update(pkg.State.from_array(scan["x_prior"]), np.eye(23) * 1e-4, max_iter=4)
    out = np.zeros(160)
    L.flivo_debug_sum(F.h, out.ctypes.data)
    t = np.concatenate([out[100:101], out[102:117]])
    print("trips", rep.trips, "ns deltas (0->2 reduce+load, ->3 finite, ->4 parts, ->5 Trows, ->6 Tcols, ->7 Aug, ->8 GJ, ->9 W, ->10 Kx, ->11 dxo, ->12 boxplus, ->13 stopjac, ...):", [int(b - a) for a, b in zip(t[:-1], t[1:])], "total", int(t[-1]-t[0]))
