"""-m gpu: the device-resident front end (flivo_raw_upload -> flivo_raw_undistort -> flivo_raw_downsample) through the
C ABI against the CPU oracle (orc_undistort: src/IMU_Processing.cpp:336-376; orc_voxel_grid: pcl::VoxelGrid as called at
src/laser_mapping.cpp:129-130).  Voxel grid: bit-exact (same float operations in the same order).  Undistortion: computed
in double and narrowed to float on both sides; device sin/cos may differ from libm in the last bit of the double, so a
float result may flip by one ulp: tolerance 4 ulp of the coordinate magnitude, and >= 99 % of the values bit-equal."""
import numpy as np
import pytest

import flivo_frontend as fe
import flivo_parity as par
import flivo_scenes as scenes
from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu


def _ctx(pkg, **kw):
    return pkg.Flivo(downsample_size=0.5, initial_points=1 << 14, **kw)


@pytest.mark.parametrize("n,leaf,sort_scan", [(50000, 0.5, True), (50000, 0.5, False), (200000, 0.2, True), (1000, 5.0, True)])
def test_voxel_grid_bit_exact(pkg, n, leaf, sort_scan):
    xyz, t, inten = fe.make_raw(11 + n, n, extent=20.0)
    xyz[7] = np.nan                                               # non-finite points are dropped
    xyz[n // 2, 1] = np.inf
    rec = fe.raw_records(xyz, t, inten)
    want, want_t = orc.voxel_grid(np.c_[xyz, inten].astype(np.float32), leaf, t_ms=t)
    with _ctx(pkg, sort_scan=sort_scan) as F:
        F.raw_upload(rec, time_col=9, intensity_col=8)
        m = F.raw_downsample(leaf)
        got = F.scan_download()
    assert m == want.shape[0] == got.shape[0]
    assert np.array_equal(got, want)


def test_voxel_grid_leaf_too_small_passes_input_through(pkg):
    xyz = np.array([[0, 0, 0], [3000, 3000, 3000], [1, 2, 3], [5, 5, 5]], np.float32)
    rec = fe.raw_records(xyz, np.zeros(4, np.float32), np.arange(4, dtype=np.float32))
    with _ctx(pkg) as F:
        F.raw_upload(rec, time_col=9, intensity_col=8)
        assert F.raw_downsample(0.001) == 4
        got = F.scan_download()
    assert np.array_equal(got[:, :3], xyz) and np.array_equal(got[:, 3], np.arange(4, dtype=np.float32))


def test_voxel_grid_single_point_and_repeat(pkg):
    rec = fe.raw_records(np.array([[1.5, -2.5, 0.25]], np.float32), np.zeros(1, np.float32), np.ones(1, np.float32))
    with _ctx(pkg) as F:
        for _ in range(2):                                        # buffers are re-used across calls
            F.raw_upload(rec, time_col=9, intensity_col=8)
            assert F.raw_downsample(0.5) == 1
            assert np.array_equal(F.scan_download()[0], np.array([1.5, -2.5, 0.25, 1.0], np.float32))


def test_undistort_matches_oracle(pkg):
    n = 100000
    xyz, t, inten = fe.make_raw(21, n, extent=60.0)
    poses, rot_end, pos_end = fe.make_poses(22)
    R_LI = scenes.exp_so3(np.array([0.01, -0.02, 0.03]))
    t_LI = np.array([0.04165, 0.02326, -0.0284])
    want = orc.undistort(xyz, t, poses, rot_end, pos_end, R_LI, t_LI)
    with _ctx(pkg) as F:
        F.raw_upload(fe.raw_records(xyz, t, inten), time_col=9, intensity_col=8)
        F.raw_undistort(poses, rot_end, pos_end, R_LI, t_LI)
        got = F.raw_download()
    assert np.array_equal(got[:, 3], inten)
    d = np.abs(got[:, :3] - want)
    assert d.max() <= 4 * np.spacing(np.float32(80.0))
    assert (got[:, :3] == want).mean() > 0.99
    assert np.array_equal(got[0, :3], xyz[0])                     # the point at time 0 is left alone


def test_undistort_needs_times_and_poses(pkg):
    xyz, t, inten = fe.make_raw(1, 100)
    poses, rot_end, pos_end = fe.make_poses(2)
    with _ctx(pkg) as F:
        with pytest.raises(RuntimeError):
            F.raw_undistort(poses, rot_end, pos_end, np.eye(3), np.zeros(3))      # nothing uploaded
        F.raw_upload(fe.raw_records(xyz, t, inten), time_col=None, intensity_col=8)
        with pytest.raises(RuntimeError):
            F.raw_undistort(poses, rot_end, pos_end, np.eye(3), np.zeros(3))      # no time stamps
        F.raw_upload(fe.raw_records(xyz, t, inten), time_col=9, intensity_col=8)
        with pytest.raises(RuntimeError):
            F.raw_undistort(poses[:1], rot_end, pos_end, np.eye(3), np.zeros(3))  # one pose is not a segment


def test_pipeline_raw_to_update_equals_host_downsample(pkg, small_world):
    """raw scan -> device voxel grid -> IEKF update  ==  oracle voxel grid on the host -> flivo_scan_upload -> IEKF update"""
    scan = scenes.make_scan(small_world, seed=77, n_points=30000, kind="avia")
    body = scan["body"].astype(np.float32)
    rng = np.random.default_rng(5)
    raw = np.repeat(body, 3, axis=0) + rng.normal(0, 0.02, (3 * len(body), 3)).astype(np.float32)
    t = np.sort(rng.uniform(0, 100, len(raw))).astype(np.float32)
    inten = rng.uniform(0, 255, len(raw)).astype(np.float32)
    leaf = 0.5
    down, _ = orc.voxel_grid(np.c_[raw, inten].astype(np.float32), leaf, t_ms=t)
    P0 = np.eye(23) * 1e-4
    cfg = pkg.make_cfg(mode=0, max_iter=4, laser_point_cov=0.001)
    res = []
    for device_side in (True, False):
        with pkg.Flivo(downsample_size=small_world["ds"], initial_points=len(small_world["points"])) as F:
            F.Build(small_world["points"])
            if device_side:
                F.raw_upload(fe.raw_records(raw, t, inten), time_col=9, intensity_col=8)
                assert F.raw_downsample(leaf) == len(down)
            else:
                F.scan_upload(np.ascontiguousarray(down[:, :3]))
            x, P, rep = F.iekf_update(pkg.State.from_array(scan["x_prior"]), P0, cfg)
            res.append((x.as_array(), np.asarray(P).copy(), rep.trips, rep.n_eff))
    assert res[0][2] == res[1][2] and res[0][3] == res[1][3]
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])
