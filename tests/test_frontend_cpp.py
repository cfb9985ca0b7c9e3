"""-m gpu: the front end through the C++ drop-in classes (host/imu_processing.h `ImuProcess::UndistortPcl`,
host/laser_mapping.h `downsample_on_device`, `lasermap_fov_segment`) driven by tests/cpp/test_dropin.cpp --frontend.
Forward propagation (src/IMU_Processing.cpp:237-334) is compared with a numpy restatement written here, the per-point
part with the oracle (orc_undistort / orc_voxel_grid), the sliding local map (src/origin_laserMapping.cpp:367-429) with a
literal Python restatement over a brute-force point set."""
import os
import struct
import subprocess

import numpy as np
import pytest

import flivo_frontend as fe
import flivo_scenes as scenes
from oracle import pyoracle as orc

pytestmark = pytest.mark.gpu
G = 9.81


def _forward_numpy(x, cov, last_imu, imu, lidar_beg, last_end, pcl_end, mean_acc,
                   cov_acc=0.1, cov_gyr=0.1, cov_bg=0.1, cov_ba=0.1):
    """src/IMU_Processing.cpp:237-334 (V1 18-state order rot,pos,vel,bg,ba,grav)."""
    R, pos, vel = x[0:9].reshape(3, 3).copy(), x[9:12].copy(), x[24:27].copy()
    bg, ba, grav = x[27:30], x[30:33], x[33:36]
    v = np.vstack([last_imu[None, :], imu])
    acc_imu, w_avr = np.zeros(3), np.zeros(3)
    poses = [np.r_[0.0, acc_imu, w_avr, vel, pos, R.reshape(-1)]]
    P = cov.copy()
    skew = lambda a: np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    for k in range(len(v) - 1):
        head, tail = v[k], v[k + 1]
        if tail[0] < last_end:
            continue
        w_avr = 0.5 * (head[1:4] + tail[1:4]) - bg
        a_avr = 0.5 * (head[4:7] + tail[4:7]) * G / np.linalg.norm(mean_acc) - ba
        dt = tail[0] - last_end if head[0] < last_end else tail[0] - head[0]
        F = np.eye(18); W = np.zeros((18, 18))
        F[0:3, 0:3] = scenes.exp_so3(-w_avr * dt)
        F[0:3, 9:12] = -np.eye(3) * dt
        F[3:6, 6:9] = np.eye(3) * dt
        F[6:9, 0:3] = -R @ skew(a_avr) * dt
        F[6:9, 12:15] = -R * dt
        F[6:9, 15:18] = np.eye(3) * dt
        W[0:3, 0:3] = np.eye(3) * cov_gyr * dt * dt
        W[6:9, 6:9] = R @ (np.eye(3) * cov_acc) @ R.T * dt * dt
        W[9:12, 9:12] = np.eye(3) * cov_bg * dt * dt
        W[12:15, 12:15] = np.eye(3) * cov_ba * dt * dt
        P = F @ P @ F.T + W
        R = R @ scenes.exp_so3(w_avr * dt)
        acc_imu = R @ a_avr + grav
        pos = pos + vel * dt + 0.5 * acc_imu * dt * dt
        vel = vel + acc_imu * dt
        poses.append(np.r_[tail[0] - lidar_beg, acc_imu, w_avr, vel, pos, R.reshape(-1)])
    imu_end = v[-1, 0]
    ref_t = imu_end if imu_end > lidar_beg else lidar_beg
    note = 1.0 if pcl_end > ref_t else -1.0
    dt = note * (pcl_end - ref_t)
    vel_end = vel + note * acc_imu * dt
    rot_end = R @ scenes.exp_so3(note * w_avr * dt)
    pos_end = pos + note * vel * dt + note * 0.5 * acc_imu * dt * dt
    return np.array(poses), rot_end, pos_end, vel_end, P


def _fov_python(points, walk, cube_len, det_range):
    """src/origin_laserMapping.cpp:367-429 with a brute-force Delete_Point_Boxes (half-open boxes, :648)."""
    MOV = np.float32(1.5)
    alive = np.ones(len(points), bool)
    lo = hi = None
    out, pending = [], 0
    for p in walk:
        boxes, deleted, cache = [], 0, None
        if lo is None:
            lo = (p - cube_len / 2.0).astype(np.float32); hi = (p + cube_len / 2.0).astype(np.float32)
            out.append((0, 0, int(alive.sum()), None, lo.copy(), hi.copy()))
            continue
        d0 = np.abs(p - lo.astype(np.float64)).astype(np.float32); d1 = np.abs(p - hi.astype(np.float64)).astype(np.float32)
        thr = MOV * np.float32(det_range)
        if not ((d0 <= thr) | (d1 <= thr)).any():
            out.append((0, 0, int(alive.sum()), None, lo.copy(), hi.copy()))
            continue
        mov = np.float32(max((cube_len - 2.0 * float(MOV) * det_range) * 0.5 * 0.9, float(np.float32(det_range) * (MOV - np.float32(1)))))
        nlo, nhi = lo.copy(), hi.copy()
        for i in range(3):
            blo, bhi = lo.copy(), hi.copy()
            if d0[i] <= thr:
                nhi[i] -= mov; nlo[i] -= mov
                blo[i] = hi[i] - mov
                boxes.append((blo, bhi))
            elif d1[i] <= thr:
                nhi[i] += mov; nlo[i] += mov
                bhi[i] = lo[i] + mov
                boxes.append((blo, bhi))
        lo, hi = nlo, nhi
        cache, pending = pending, 0
        for blo, bhi in boxes:
            inside = alive & (points >= blo).all(1) & (points < bhi).all(1)
            deleted += int(inside.sum())
            alive &= ~inside
        pending += deleted
        out.append((len(boxes), deleted, int(alive.sum()), cache, lo.copy(), hi.copy()))
    return out


def test_frontend_through_dropin_classes(pkg, tmp_path):
    import importlib
    exe = importlib.import_module("fast_livo_noted_b200.build").build_host_test()
    rng = np.random.default_rng(31)
    n_raw, n_imu = 40000, 21
    xyz, t, inten = fe.make_raw(32, n_raw, extent=40.0)
    lidar_beg = 100.0
    x = np.zeros(36)
    x[0:9] = scenes.exp_so3(rng.uniform(-0.1, 0.1, 3)).reshape(-1)
    x[9:12] = rng.uniform(-2, 2, 3)
    x[12:21] = scenes.exp_so3(np.array([0.01, -0.02, 0.015])).reshape(-1)
    x[21:24] = scenes.T_LI
    x[24:27] = rng.uniform(-2, 2, 3)
    x[27:30] = rng.uniform(-0.01, 0.01, 3); x[30:33] = rng.uniform(-0.05, 0.05, 3)
    x[33:36] = np.array([0.0, 0.0, -9.81])
    stamps = lidar_beg - 0.002 + 0.005 * np.arange(n_imu + 1)          # 200 Hz, first sample = last_imu_ of the previous frame
    imu = np.zeros((n_imu + 1, 7))
    imu[:, 0] = stamps
    imu[:, 1:4] = rng.uniform(-0.5, 0.5, (n_imu + 1, 3))
    imu[:, 4:7] = np.array([0.0, 0.0, 9.81]) + rng.uniform(-1.0, 1.0, (n_imu + 1, 3))
    n_map, n_walk = 60000, 12
    cube_len, det_range = 200.0, 40.0
    cloud = rng.uniform(-250, 250, (n_map, 3)).astype(np.float32)
    walk = np.zeros((n_walk, 3)); walk[:, 0] = np.linspace(0, 260, n_walk); walk[:, 1] = np.linspace(0, -90, n_walk)
    raw5 = np.c_[xyz, inten, t].astype(np.float32)
    fin, fout = os.path.join(tmp_path, "fe_in.bin"), os.path.join(tmp_path, "fe_out.bin")
    with open(fin, "wb") as f:
        f.write(struct.pack("4i", n_raw, n_imu, n_map, n_walk))
        f.write(struct.pack("5d", 0.5, 0.5, cube_len, det_range, lidar_beg))
        f.write(x.tobytes()); f.write(imu.tobytes()); f.write(raw5.tobytes()); f.write(cloud.tobytes()); f.write(walk.tobytes())
    r = subprocess.run([exe, "--frontend", fin, fout], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    buf = open(fout, "rb").read()
    npz = struct.unpack_from("i", buf, 0)[0]; off = 4
    poses = np.frombuffer(buf, np.float64, npz * 22, off).reshape(npz, 22); off += npz * 22 * 8
    st = np.frombuffer(buf, np.float64, 15, off); off += 15 * 8
    cov = np.frombuffer(buf, np.float64, 324, off).reshape(18, 18); off += 324 * 8
    und = np.frombuffer(buf, np.float32, n_raw * 4, off).reshape(n_raw, 4); off += n_raw * 16
    nd = struct.unpack_from("i", buf, off)[0]; off += 4
    down = np.frombuffer(buf, np.float32, nd * 4, off).reshape(nd, 4)

    # forward propagation vs numpy
    pcl_end = lidar_beg + float(t[-1]) / 1000.0
    wp, rot_end, pos_end, vel_end, Pw = _forward_numpy(x, np.eye(18) * 0.001, imu[0], imu[1:], lidar_beg, lidar_beg, pcl_end, x[33:36])
    assert poses.shape == wp.shape
    assert np.abs(poses - wp).max() < 1e-9
    assert np.abs(st[0:9] - rot_end.reshape(-1)).max() < 1e-12 and np.abs(st[9:12] - pos_end).max() < 1e-12 and np.abs(st[12:15] - vel_end).max() < 1e-12
    assert np.abs(cov - Pw).max() < 1e-12
    # per-point part vs the oracle, fed with the C++ side's own poses and end state
    want = orc.undistort(xyz, t, poses, st[0:9].reshape(3, 3), st[9:12], x[12:21].reshape(3, 3), x[21:24])
    assert np.abs(und[:, :3] - want).max() <= 4 * np.spacing(np.float32(64.0))
    assert (und[:, :3] == want).mean() > 0.99 and np.array_equal(und[:, 3], inten)
    # voxel grid of what the device holds
    wd, _ = orc.voxel_grid(und, 0.5, t_ms=t)
    assert nd == len(wd) and np.array_equal(down, wd)
    lines = {l.split()[0]: l.split()[1:] for l in r.stdout.strip().splitlines() if not l.startswith("fov")}
    assert lines["run"][1] == "0" and int(lines["run"][3]) > 0 and lines["run"][5] == "1"      # first frame built the map
    # ... from feats_down_body as it is (lidar frame, no down-sampling, non-identity extrinsics ignored), as
    # src/laser_mapping.cpp:134-142 does: one map point per scan point
    assert int(lines["run"][3]) == nd
    # sliding local map
    fov = [l.split() for l in r.stdout.strip().splitlines() if l.startswith("fov")]
    exp = _fov_python(cloud, walk, cube_len, det_range)
    assert len(fov) == len(exp)
    moved = 0
    for got, (nb, dele, valid, cache, lo, hi) in zip(fov, exp):
        assert int(got[3]) == nb and int(got[5]) == dele and int(got[7]) == valid
        if cache is not None:
            assert int(got[9]) == cache
        assert np.array_equal(np.array(got[11:14], np.float32), lo) and np.array_equal(np.array(got[14:17], np.float32), hi)
        moved += nb > 0
    assert moved >= 2 and exp[-1][2] < n_map
