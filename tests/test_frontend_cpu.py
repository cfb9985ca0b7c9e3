"""CPU: the oracle's restatements of the two steps before the path (ImuProcess::UndistortPcl back-propagation,
src/IMU_Processing.cpp:336-376; pcl::VoxelGrid::filter, src/laser_mapping.cpp:129-130) against independent numpy float64
statements and analytic cases.  The reference holds no tests or fixtures for either: parity unpinned."""
import numpy as np

import flivo_frontend as fe
import flivo_scenes as scenes
from oracle import pyoracle as orc


def test_voxel_grid_centroids_and_order():
    xyz, t, inten = fe.make_raw(3, 20000, extent=12.0)
    xyzi = np.c_[xyz, inten].astype(np.float32)
    out, out_t = orc.voxel_grid(xyzi, 0.5, t_ms=t)
    want, keys = fe.voxel_grid_numpy(xyzi, 0.5)
    assert out.shape[0] == want.shape[0] and np.all(np.diff(keys) > 0)      # one output per leaf, ascending leaf index
    assert np.abs(out - want).max() < 2e-4                                   # float32 sums vs float64 means
    assert out_t.shape[0] == out.shape[0] and out_t.min() >= 0 and out_t.max() <= 100.0


def test_voxel_grid_single_point_leaves_are_bit_exact_and_nan_dropped():
    rng = np.random.default_rng(0)
    g = np.stack(np.meshgrid(np.arange(10), np.arange(10), np.arange(3), indexing="ij"), -1).reshape(-1, 3)
    xyz = (g + rng.uniform(0.1, 0.9, g.shape)).astype(np.float32)            # one point per unit leaf
    xyzi = np.c_[xyz, np.arange(len(xyz))].astype(np.float32)
    bad = xyzi[:5].copy(); bad[:, 0] = np.nan
    perm = rng.permutation(len(xyzi) + 5)
    mixed = np.r_[xyzi, bad][perm]
    out, _ = orc.voxel_grid(mixed, 1.0)
    assert out.shape[0] == len(xyzi)
    # leaf order = x fastest, then y, then z (divb_mul = (1, dx, dx*dy))
    order = np.lexsort((g[:, 0], g[:, 1], g[:, 2]))
    assert np.array_equal(out, xyzi[order])


def test_voxel_grid_leaf_too_small_returns_input():
    xyz = np.array([[0, 0, 0], [3000, 3000, 3000], [1, 2, 3]], np.float32)
    xyzi = np.c_[xyz, [1, 2, 3]].astype(np.float32)
    out, _ = orc.voxel_grid(xyzi, 0.001)          # 3e6^3 leaves > INT_MAX: PCL warns and copies the input
    assert np.array_equal(out, xyzi)


def test_undistort_matches_numpy_statement():
    xyz, t, _ = fe.make_raw(5, 5000)
    poses, rot_end, pos_end = fe.make_poses(6)
    R_LI = scenes.exp_so3(np.array([0.01, -0.02, 0.03]))
    t_LI = np.array([0.04165, 0.02326, -0.0284])
    got = orc.undistort(xyz, t, poses, rot_end, pos_end, R_LI, t_LI)
    want = fe.undistort_numpy(xyz, t, poses, rot_end, pos_end, R_LI, t_LI)
    assert np.abs(got - want).max() < 1e-4                                   # float32 output, |coordinates| <= 30
    assert np.array_equal(got[0], xyz[0])                                    # t = 0: never reached by the loop (:354)
    assert np.abs(got - xyz).max() > 0.05                                    # the motion is not a no-op


def test_undistort_pure_translation_is_analytic():
    # no rotation, constant velocity v, identity extrinsic: p_end = p + (pos(t) - pos_end)
    n = 1000
    rng = np.random.default_rng(2)
    xyz = rng.uniform(-20, 20, (n, 3)).astype(np.float32)
    t = np.linspace(0, 100, n).astype(np.float32)
    v = np.array([2.0, -1.0, 0.5])
    poses = np.zeros((11, 22))
    for k in range(11):
        poses[k, 0] = 0.01 * k
        poses[k, 7:10] = v
        poses[k, 10:13] = v * 0.01 * k
        poses[k, 13:22] = np.eye(3).reshape(-1)
    pos_end = v * 0.1
    got = orc.undistort(xyz, t, poses, np.eye(3), pos_end, np.eye(3), np.zeros(3))
    want = xyz.astype(np.float64) + (v[None, :] * (t[:, None].astype(np.float64) / 1000.0) - pos_end[None, :])
    want[0] = xyz[0]
    assert np.abs(got - want).max() < 5e-6
