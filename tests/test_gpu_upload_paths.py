"""-m gpu: the three ways a scan of 48-byte PointXYZINormal-like records reaches the GPU through flivo_scan_upload_async
(csrc/api.cu) must give the same update, bit for bit: pageable memory (host repack to float4), registered memory with the
repack (many host threads), registered memory sent as it is (FLIVO_RAW_DMA=1), and the split a rank with few host threads
uses (FLIVO_RAW_DMA=2: head of the scan repacked by the host, tail sent raw).
FLIVO_RAW_DMA is read once per process, so every flavour runs in a process of its own and prints its posterior as hex."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(flavour):
    env = dict(os.environ)
    env.pop("FLIVO_RAW_DMA", None)
    if flavour == "raw":
        env["FLIVO_RAW_DMA"] = "1"
    elif flavour == "registered_repack":
        env["FLIVO_RAW_DMA"] = "0"
    elif flavour == "split":                       # head repacked by the host, tail raw: what a rank with few host threads does
        env["FLIVO_RAW_DMA"] = "2"
        env["FLIVO_RAW_SPLIT"] = "37"
    r = subprocess.run([sys.executable, os.path.abspath(__file__), flavour], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1].split()
    return line[1:]


def test_upload_flavours_give_the_same_update():
    ref = _run("pageable")
    assert int(ref[0]) > 1000                      # n_eff: the update saw the scan
    for flavour in ("registered_repack", "raw", "split"):
        got = _run(flavour)
        assert got == ref, flavour


if __name__ == "__main__":
    flavour = sys.argv[1]
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ctypes as C
    import __graft_entry__ as g
    import flivo_scenes as scenes
    pkg = g.load_package()
    world = scenes.make_world(seed=11, n_points=60000, ds=0.5)
    n = 70000                                      # above the size from which a rank with few host threads sends raw records
    scan = scenes.make_scan(world, seed=3, n_points=n, pos=scenes.street_positions(world, 1)[0])
    rec = np.zeros((n, 12), np.float32)
    rec[:, :3] = scan["body"]
    rec[:, 3:] = np.random.default_rng(1).uniform(-1, 1, (n, 9)).astype(np.float32)   # intensity, normals, curvature: must be ignored
    L = pkg.lib()
    if flavour != "pageable":
        assert L.flivo_host_register(C.c_void_p(rec.ctypes.data), rec.nbytes) == 0
    with pkg.Flivo(downsample_size=world["ds"], initial_points=len(world["points"]), sort_scan=True) as F:
        F.Build(world["points"])
        out = []
        for _ in range(2):                          # both landing slots
            F.scan_upload_async(rec)
            F.scan_commit()
            x, P, rep = F.iekf_update(pkg.State.from_array(scan["x_prior"]), np.eye(23) * 1e-4, mode=0, max_iter=4)
            out.append((rep.n_eff, x.as_array().tobytes().hex(), np.asarray(P).tobytes().hex()))
        assert out[0] == out[1]
    if flavour != "pageable":
        L.flivo_host_unregister(C.c_void_p(rec.ctypes.data))
    print("RESULT", out[0][0], out[0][1], out[0][2][:256])
