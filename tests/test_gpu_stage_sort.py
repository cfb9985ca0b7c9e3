"""-m gpu: the staging sort of a scan (csrc/stage.cu) against numpy.  The processing order must be the STABLE sort of the
scan by bits [6, 22) of the Morton code of each point's search cell in the lidar frame — equal keys keep the caller's
order, which is what makes every fixed-order sum downstream a pure function of the scan.  Sizes cover a partial tile,
several tiles, a padded tail, the 8192-point-tile flavour (> 2^20 points) and heavy ties (all points in a few cells)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _spread21(v):
    x = v.astype(np.uint64) & np.uint64(0x1FFFFF)
    x = (x | (x << np.uint64(32))) & np.uint64(0x1F00000000FFFF)
    x = (x | (x << np.uint64(16))) & np.uint64(0x1F0000FF0000FF)
    x = (x | (x << np.uint64(8))) & np.uint64(0x100F00F00F00F00F)
    x = (x | (x << np.uint64(4))) & np.uint64(0x10C30C30C30C30C3)
    x = (x | (x << np.uint64(2))) & np.uint64(0x1249249249249249)
    return x


def _key16(xyz, cell_edge):
    inv = np.float32(1.0) / np.float32(cell_edge)
    c = np.clip(np.floor(xyz.astype(np.float32) * inv), -1048575.0, 1048575.0).astype(np.int64) + (1 << 20)
    m = (_spread21(c[:, 0]) << np.uint64(2)) | (_spread21(c[:, 1]) << np.uint64(1)) | _spread21(c[:, 2])
    return ((m >> np.uint64(6)) & np.uint64(0xFFFF)).astype(np.int64)


def _scan(kind, n, rng):
    if kind == "room":          # ties everywhere: a few hundred cells
        return rng.uniform(-3.0, 3.0, (n, 3)).astype(np.float32)
    if kind == "street":
        xyz = rng.uniform(-60.0, 60.0, (n, 3)).astype(np.float32)
        xyz[:, 2] = rng.uniform(-2.0, 12.0, n).astype(np.float32)
        return xyz
    if kind == "one_cell":
        return (rng.uniform(0.01, 0.39, (n, 3)) + np.array([4.0, -8.0, 0.4])).astype(np.float32)
    raise ValueError(kind)


def _check_order(pkg, kind, n):
    rng = np.random.default_rng(1000 + n)
    ds, ratio = 0.2, 2
    xyz = _scan(kind, n, rng)
    with pkg.Flivo(downsample_size=ds, cell_ratio=ratio, initial_points=1000, sort_scan=True) as F:
        F.Build(rng.uniform(-5, 5, (1000, 3)).astype(np.float32))
        F.scan_upload(xyz)
        got = F.scan_order(n).astype(np.int64)
        want = np.argsort(_key16(xyz, np.float32(ds) * np.float32(ratio)), kind="stable")
        assert np.array_equal(np.sort(got), np.arange(n)), "not a permutation"
        assert np.array_equal(got, want)
        # a second, different scan of another size on the same context (tickets and histograms must come back clean)
        m = max(1, n // 3 + 5)
        xyz2 = _scan("street", m, rng)
        F.scan_upload(xyz2)
        got2 = F.scan_order(m).astype(np.int64)
        assert np.array_equal(got2, np.argsort(_key16(xyz2, np.float32(ds) * np.float32(ratio)), kind="stable"))
        # and the gathered scan is the caller's scan in that order
        body = F.scan_download()
        assert np.array_equal(body[:, :3], xyz2)


CASES = [("street", 1), ("street", 777), ("room", 2048), ("street", 24000), ("room", 60000), ("one_cell", 5000),
         ("room", 100000), ("street", 260000), ("street", 1200000)]


@pytest.mark.parametrize("kind,n", CASES)
def test_processing_order_is_the_stable_sort_by_cell_key(pkg, kind, n):
    """Default choice of implementation: the own sort up to 2^20 points, the library's above."""
    _check_order(pkg, kind, n)


@pytest.mark.parametrize("which", ["own", "cub"])
def test_both_implementations_at_every_size(which):
    """FLIVO_SORT is read once per process, so each forced flavour runs in a process of its own (own: incl. the 8192-point
    tiles above 2^20 points; cub: the small sizes the default never hands it)."""
    env = dict(os.environ, FLIVO_SORT=which)
    r = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "checked %d cases" % len(CASES) in r.stdout


def test_records_with_a_stride_are_staged_in_place(pkg):
    """48-byte PointXYZINormal-like records (x y z first) through the async upload path give the same order."""
    rng = np.random.default_rng(5)
    n, ds, ratio = 30000, 0.5, 2
    rec = np.zeros((n, 12), np.float32)
    rec[:, :3] = _scan("street", n, rng)
    rec[:, 3:] = rng.uniform(-1, 1, (n, 9)).astype(np.float32)
    with pkg.Flivo(downsample_size=ds, cell_ratio=ratio, initial_points=1000, sort_scan=True) as F:
        F.Build(rng.uniform(-5, 5, (1000, 3)).astype(np.float32))
        F.scan_upload_async(rec)
        F.scan_commit()
        got = F.scan_order(n).astype(np.int64)
        assert np.array_equal(got, np.argsort(_key16(rec[:, :3], np.float32(ds) * np.float32(ratio)), kind="stable"))


if __name__ == "__main__":
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    _pkg = g.load_package()
    for _kind, _n in CASES:
        _check_order(_pkg, _kind, _n)
    print("checked %d cases with FLIVO_SORT=%s" % (len(CASES), os.environ.get("FLIVO_SORT", "(default)")))
