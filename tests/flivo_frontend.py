"""Helpers of the front-end tests: synthetic raw scans with time stamps, IMU pose chains, and
independent numpy float64 statements of the two steps (used to check the C++ oracle itself)."""
import numpy as np

import flivo_scenes as scenes


def raw_records(xyz, t_ms, intensity):
    """(n, 12) float32 records in the reference's pcl::PointXYZINormal layout: xyz at floats 0..2, intensity at 8,
    curvature (time offset in ms) at 9."""
    rec = np.zeros((len(xyz), 12), np.float32)
    rec[:, :3] = xyz
    rec[:, 8] = intensity
    rec[:, 9] = t_ms
    return rec


def make_raw(seed, n, extent=30.0, scan_ms=100.0):
    rng = np.random.default_rng(seed)
    xyz = rng.uniform(-extent, extent, (n, 3)).astype(np.float32)
    xyz[:, 2] *= 0.2
    t = np.sort(rng.uniform(0.0, scan_ms, n)).astype(np.float32)
    t[0] = 0.0                                   # the reference's first point has offset 0 and is never compensated
    inten = rng.uniform(0, 255, n).astype(np.float32)
    return xyz, t, inten


def make_poses(seed, n_poses=21, scan_s=0.1):
    """IMUpose as ImuProcess::UndistortPcl builds it: constant-acceleration, constant-rate segments, pose k at k*dt.
    Returns (poses (m,22), rot_end, pos_end)."""
    rng = np.random.default_rng(seed)
    dt = scan_s / (n_poses - 1)
    R = scenes.exp_so3(rng.uniform(-0.2, 0.2, 3))
    pos = rng.uniform(-5, 5, 3)
    vel = rng.uniform(-3, 3, 3)
    poses = np.zeros((n_poses, 22))
    gyr = rng.uniform(-1.0, 1.0, 3)
    acc = rng.uniform(-2, 2, 3)
    for k in range(n_poses):
        poses[k, 0] = k * dt
        poses[k, 1:4] = acc
        poses[k, 4:7] = gyr
        poses[k, 7:10] = vel
        poses[k, 10:13] = pos
        poses[k, 13:22] = R.reshape(-1)
        # propagate to the next sample (src/IMU_Processing.cpp:300-318)
        R = R @ scenes.exp_so3(gyr * dt)
        pos = pos + vel * dt + 0.5 * acc * dt * dt
        vel = vel + acc * dt
        gyr = gyr + rng.uniform(-0.1, 0.1, 3)
        acc = acc + rng.uniform(-0.3, 0.3, 3)
    # the frame-end pose: a little after the last IMU sample (:321-330)
    rot_end = poses[-1, 13:22].reshape(3, 3) @ scenes.exp_so3(poses[-1, 4:7] * 0.002)
    pos_end = poses[-1, 10:13] + poses[-1, 7:10] * 0.002
    return poses, rot_end, pos_end


def undistort_numpy(xyz, t_ms, poses, rot_end, pos_end, R_LI, t_LI):
    """float64 statement of the per-point rule: head = last pose (excluding the final one) whose offset is < t."""
    out = xyz.astype(np.float64).copy()
    t = t_ms.astype(np.float64) / 1000.0
    off = poses[:-1, 0]
    head = np.searchsorted(off, t, side="left") - 1          # largest k with off[k] < t
    extR_Ri = R_LI.T @ rot_end.T
    ext_t = R_LI.T @ t_LI
    for i in np.nonzero(head >= 0)[0]:
        h = poses[head[i]]
        dt = t[i] - h[0]
        R_i = h[13:22].reshape(3, 3) @ scenes.exp_so3(h[4:7] * dt)
        T_ei = h[10:13] + h[7:10] * dt + 0.5 * h[1:4] * dt * dt - pos_end
        out[i] = extR_Ri @ (R_i @ (R_LI @ out[i] + t_LI) + T_ei) - ext_t
    return out


def voxel_grid_numpy(xyzi, leaf):
    """float64 centroids per leaf, ordered by leaf index (keys computed in float32 like PCL)."""
    x = xyzi[np.isfinite(xyzi[:, :3]).all(1)]
    inv = np.float32(1.0) / np.float32(leaf)
    mnb = np.floor(x[:, :3].min(0) * inv).astype(np.int64)
    mxb = np.floor(x[:, :3].max(0) * inv).astype(np.int64)
    div = mxb - mnb + 1
    ijk = (np.floor(x[:, :3] * inv) - mnb.astype(np.float32)).astype(np.int64)
    idx = ijk[:, 0] + ijk[:, 1] * div[0] + ijk[:, 2] * div[0] * div[1]
    order = np.argsort(idx, kind="stable")
    idx_s = idx[order]
    starts = np.r_[0, np.nonzero(np.diff(idx_s))[0] + 1]
    sums = np.add.reduceat(x[order].astype(np.float64), starts, axis=0)
    cnt = np.diff(np.r_[starts, len(idx_s)])
    return sums / cnt[:, None], idx_s[starts]
