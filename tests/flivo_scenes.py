"""Deterministic synthetic worlds and scans shared by tests/, smoke() and bench.py.

World: "city block" = ground plane + axis-aligned boxes; points sampled on the surfaces with N(0, 0.01 m)
normal noise and reduced by the Add_Points voxel rule (one point per downsample voxel, the one nearest the
voxel centre; reference include/ikd-Tree/ikd_Tree.cpp:392-417), to exactly n_points.
Scan: analytic ray casting of an Avia-like (70.4 x 77.2 deg) or Ouster-128-like (360 x 45 deg) pattern
against the same geometry, N(0, 0.02 m) range noise, range in [1, 120] m, expressed in the lidar frame.
Pure numpy; nothing here touches the GPU, the product library or /root/reference.
"""
import numpy as np

T_LI = np.array([0.04165, 0.02326, -0.0284])   # lidar -> IMU offset (Avia default style)


def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def exp_so3(v):
    th = np.linalg.norm(v)
    K = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])
    if th < 1e-12:
        return np.eye(3) + K
    return np.eye(3) + np.sin(th) / th * K + (1 - np.cos(th)) / th ** 2 * (K @ K)


def voxel_downsample(pts, ds):
    """One point per voxel: nearest to the voxel centre, first on ties (float32, reference formulas)."""
    ds32 = np.float32(ds)
    k = np.floor(pts / ds32).astype(np.int64)
    lo = (k.astype(np.float32) * ds32)
    mid = (lo.astype(np.float64) + (((lo + ds32) - lo).astype(np.float64)) / 2.0).astype(np.float32)
    d = pts - mid
    dist = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
    key = ((k[:, 0] + (1 << 20)) << 42) | ((k[:, 1] + (1 << 20)) << 21) | (k[:, 2] + (1 << 20))
    order = np.lexsort((np.arange(len(pts)), dist, key))
    ks = key[order]
    head = np.ones(len(ks), bool)
    head[1:] = ks[1:] != ks[:-1]
    return pts[order[head]]


def make_world(seed, n_points, ds, pitch=30.0):
    """-> dict(points (n,3) f32, ds, extent, boxes (B,5) = cx, cy, w, d, h)."""
    rng = np.random.default_rng(0xF4570000 + seed)
    area_needed = n_points * ds * ds
    extent = max(float(np.sqrt(0.42 * area_needed)), 2.2 * pitch)
    nb = max(int(extent // pitch), 1)
    cx, cy = np.meshgrid((np.arange(nb) + 0.5) * pitch - nb * pitch / 2, (np.arange(nb) + 0.5) * pitch - nb * pitch / 2)
    B = cx.size
    boxes = np.stack([cx.ravel() + rng.uniform(-3, 3, B), cy.ravel() + rng.uniform(-3, 3, B),
                      rng.uniform(8, 20, B), rng.uniform(8, 20, B), rng.uniform(3, 20, B)], axis=1)
    # keep the block that holds the origin free so a sensor can stand there
    boxes = boxes[np.hypot(boxes[:, 0], boxes[:, 1]) > 16.0]
    # rectangles: origin o, edges u, v, unit normal nrm
    rects = [(np.array([-extent / 2, -extent / 2, 0.0]), np.array([extent, 0, 0.0]), np.array([0, extent, 0.0]))]
    for bx, by, w, d, h in boxes:
        x0, x1, y0, y1 = bx - w / 2, bx + w / 2, by - d / 2, by + d / 2
        rects += [(np.array([x0, y0, 0.0]), np.array([w, 0, 0.0]), np.array([0, 0, h])),
                  (np.array([x0, y1, 0.0]), np.array([w, 0, 0.0]), np.array([0, 0, h])),
                  (np.array([x0, y0, 0.0]), np.array([0, d, 0.0]), np.array([0, 0, h])),
                  (np.array([x1, y0, 0.0]), np.array([0, d, 0.0]), np.array([0, 0, h])),
                  (np.array([x0, y0, h]), np.array([w, 0, 0.0]), np.array([0, d, 0.0]))]
    O = np.array([r[0] for r in rects]); U = np.array([r[1] for r in rects]); V = np.array([r[2] for r in rects])
    Nn = np.cross(U, V)
    areas = np.linalg.norm(Nn, axis=1)
    Nn = Nn / areas[:, None]
    out = np.zeros((0, 3), np.float32)
    factor = 1.7
    for _ in range(6):
        n_cand = int(factor * areas.sum() / (ds * ds)) + 1000
        chunks = []
        left = n_cand
        while left > 0:                       # chunked to bound memory at 20M+ points
            m = min(left, 8_000_000)
            ridx = rng.choice(len(rects), size=m, p=areas / areas.sum())
            a, b = rng.random(m), rng.random(m)
            p = O[ridx] + a[:, None] * U[ridx] + b[:, None] * V[ridx] + rng.normal(0, 0.01, m)[:, None] * Nn[ridx]
            chunks.append(voxel_downsample(p.astype(np.float32), ds))
            left -= m
        out = voxel_downsample(np.concatenate([out] + chunks), ds)
        if len(out) >= n_points:
            break
        factor *= 1.5
    if len(out) < n_points:
        raise RuntimeError("world generator could not reach %d points (got %d)" % (n_points, len(out)))
    keep = rng.permutation(len(out))[:n_points]
    keep.sort()
    pts = np.ascontiguousarray(out[keep])
    return dict(points=pts, ds=float(ds), extent=extent, boxes=boxes, seed=seed)


def _cast(origin, dirs, world, rmax=120.0):
    """Nearest hit range of rays against ground z=0 and the boxes (inf = miss)."""
    n = len(dirs)
    t = np.full(n, np.inf)
    dz = dirs[:, 2]
    with np.errstate(divide="ignore", invalid="ignore"):
        tg = np.where(dz < -1e-9, -origin[2] / dz, np.inf)
    half = world["extent"] / 2
    hit = origin[None, :2] + tg[:, None] * dirs[:, :2]
    ok = np.isfinite(tg) & (np.abs(hit[:, 0]) < half) & (np.abs(hit[:, 1]) < half)
    t = np.where(ok, tg, t)
    boxes = world["boxes"]
    near = boxes[np.hypot(boxes[:, 0] - origin[0], boxes[:, 1] - origin[1]) < rmax + 30.0]
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / dirs
    for bx, by, w, d, h in near:
        lo = np.array([bx - w / 2, by - d / 2, 0.0]); hi = np.array([bx + w / 2, by + d / 2, h])
        t0 = (lo - origin) * inv; t1 = (hi - origin) * inv
        tn = np.nanmax(np.minimum(t0, t1), axis=1); tf = np.nanmin(np.maximum(t0, t1), axis=1)
        ok = (tn <= tf) & (tn > 0)
        t = np.where(ok & (tn < t), tn, t)
    return t


def make_scan(world, seed, n_points, kind="avia", pos=None, prior_noise=True):
    """-> dict(body (n,3) f32 lidar frame, x_true (36,), x_prior (36,), P (23,23), kind)."""
    rng = np.random.default_rng(0xF4571000 + seed)
    R_true = _rot(rng.uniform(-0.03, 0.03), rng.uniform(-0.03, 0.03), rng.uniform(-np.pi, np.pi))
    t_true = np.array([rng.uniform(-4, 4), rng.uniform(-4, 4), rng.uniform(1.5, 2.2)]) if pos is None else np.asarray(pos, float)
    origin = R_true @ T_LI + t_true
    body = np.zeros((0, 3))
    tries = 0
    while len(body) < n_points:
        m = int((n_points - len(body)) * (1.6 if kind == "avia" else 2.6)) + 256
        if kind == "avia":      # forward-looking cone, rosette-like: dense centre
            az = np.deg2rad(rng.uniform(-35.2, 35.2, m)); el = np.deg2rad(rng.uniform(-38.6, 38.6, m))
            order = np.lexsort((az, np.floor(el * 40)))
            az, el = az[order], el[order]
        else:                   # Ouster-128-like: 128 rings x columns, column-major like the driver
            cols = m // 128 + 1
            azc = (np.arange(cols) + rng.random()) * (2 * np.pi / cols)
            elr = np.deg2rad(np.linspace(-22.5, 22.5, 128))
            az = np.repeat(azc, 128)[:m]; el = np.tile(elr, cols)[:m]
            az = az + rng.normal(0, 1e-4, m); el = el + rng.normal(0, 1e-4, m)
        d_b = np.stack([np.cos(el) * np.cos(az), np.cos(el) * np.sin(az), np.sin(el)], axis=1)
        r = _cast(origin, d_b @ R_true.T, world)
        ok = np.isfinite(r) & (r >= 1.0) & (r <= 120.0)
        rr = r[ok] + rng.normal(0, 0.02, ok.sum())
        body = np.concatenate([body, d_b[ok] * rr[:, None]])
        tries += 1
        if tries > 20:
            raise RuntimeError("scan generator starved")
    body = np.ascontiguousarray(body[:n_points].astype(np.float32))
    x_true = np.zeros(36)
    x_true[0:9] = R_true.reshape(9); x_true[9:12] = t_true
    x_true[12:21] = np.eye(3).reshape(9); x_true[21:24] = T_LI
    x_true[33:36] = [0.0, 0.0, -9.809]
    x_prior = x_true.copy()
    if prior_noise:
        dth = np.deg2rad(rng.uniform(-0.5, 0.5, 3)); dt = rng.uniform(-0.05, 0.05, 3)
        x_prior[0:9] = (R_true @ exp_so3(dth)).reshape(9)
        x_prior[9:12] = t_true + dt
    return dict(body=body, x_true=x_true, x_prior=x_prior, P=np.eye(23) * 1e-4, kind=kind, seed=seed)


# ------------------------------------------------------------------------------------------------------
# torch versions of the same generators for bench-size workloads (20M-point maps, 260k-point scans): the
# numpy path above takes minutes there.  Same geometry and rules; a different (torch) random stream.
# `device` may be "cuda" (bench on the GPU box) or "cpu".
# ------------------------------------------------------------------------------------------------------
def _voxel_downsample_t(pts, ds):
    import torch
    ds32 = torch.tensor(ds, dtype=torch.float32, device=pts.device)
    kf = torch.floor(pts / ds32)
    k = kf.to(torch.int64)
    lo = kf * ds32
    mid = (lo.double() + ((lo + ds32) - lo).double() / 2.0).float()
    d = pts - mid
    dist = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
    key = ((k[:, 0] + (1 << 20)) << 42) | ((k[:, 1] + (1 << 20)) << 21) | (k[:, 2] + (1 << 20))
    o1 = torch.sort(dist, stable=True).indices
    o2 = torch.sort(key[o1], stable=True)
    order = o1[o2.indices]
    ks = o2.values
    head = torch.ones(len(ks), dtype=torch.bool, device=pts.device)
    head[1:] = ks[1:] != ks[:-1]
    return pts[order[head]]


def make_world_torch(seed, n_points, ds, device="cuda", pitch=30.0):
    """make_world at bench sizes.  -> dict(points: torch (n,3) f32 on `device`, ds, extent, boxes (numpy))."""
    import torch
    rng = np.random.default_rng(0xF4570000 + seed)
    g = torch.Generator(device=device)
    g.manual_seed(0xF4570000 + seed)
    area_needed = n_points * ds * ds
    extent = max(float(np.sqrt(0.42 * area_needed)), 2.2 * pitch)
    nb = max(int(extent // pitch), 1)
    cx, cy = np.meshgrid((np.arange(nb) + 0.5) * pitch - nb * pitch / 2, (np.arange(nb) + 0.5) * pitch - nb * pitch / 2)
    B = cx.size
    boxes = np.stack([cx.ravel() + rng.uniform(-3, 3, B), cy.ravel() + rng.uniform(-3, 3, B),
                      rng.uniform(8, 20, B), rng.uniform(8, 20, B), rng.uniform(3, 20, B)], axis=1)
    boxes = boxes[np.hypot(boxes[:, 0], boxes[:, 1]) > 16.0]
    O = [np.array([-extent / 2, -extent / 2, 0.0])]; U = [np.array([extent, 0, 0.0])]; V = [np.array([0, extent, 0.0])]
    for bx, by, w, d, h in boxes:
        x0, x1, y0, y1 = bx - w / 2, bx + w / 2, by - d / 2, by + d / 2
        O += [np.array([x0, y0, 0.0]), np.array([x0, y1, 0.0]), np.array([x0, y0, 0.0]), np.array([x1, y0, 0.0]), np.array([x0, y0, h])]
        U += [np.array([w, 0, 0.0]), np.array([w, 0, 0.0]), np.array([0, d, 0.0]), np.array([0, d, 0.0]), np.array([w, 0, 0.0])]
        V += [np.array([0, 0, h]), np.array([0, 0, h]), np.array([0, 0, h]), np.array([0, 0, h]), np.array([0, d, 0.0])]
    O, U, V = np.array(O), np.array(U), np.array(V)
    Nn = np.cross(U, V)
    areas = np.linalg.norm(Nn, axis=1)
    Nn = Nn / areas[:, None]
    tO, tU, tV, tN = (torch.tensor(a, dtype=torch.float64, device=device) for a in (O, U, V, Nn))
    cdf = torch.tensor(np.cumsum(areas / areas.sum()), dtype=torch.float64, device=device)
    out = torch.zeros((0, 3), dtype=torch.float32, device=device)
    factor = 1.7
    for _ in range(6):
        left = int(factor * areas.sum() / (ds * ds)) + 1000
        chunks = [out]
        while left > 0:
            m = min(left, 16_000_000)
            ridx = torch.searchsorted(cdf, torch.rand(m, generator=g, device=device, dtype=torch.float64)).clamp_(max=len(areas) - 1)
            a = torch.rand(m, generator=g, device=device, dtype=torch.float64)
            b = torch.rand(m, generator=g, device=device, dtype=torch.float64)
            nz = torch.randn(m, generator=g, device=device, dtype=torch.float64) * 0.01
            p = tO[ridx] + a[:, None] * tU[ridx] + b[:, None] * tV[ridx] + nz[:, None] * tN[ridx]
            chunks.append(_voxel_downsample_t(p.float(), ds))
            left -= m
        out = _voxel_downsample_t(torch.cat(chunks), ds)
        if len(out) >= n_points:
            break
        factor *= 1.5
    if len(out) < n_points:
        raise RuntimeError("world generator could not reach %d points (got %d)" % (n_points, len(out)))
    keep = torch.randperm(len(out), generator=g, device=device)[:n_points].sort().values
    return dict(points=out[keep].contiguous(), ds=float(ds), extent=extent, boxes=boxes, seed=seed)


def _cast_t(origin, dirs, world, rmax=120.0):
    import torch
    dev = dirs.device
    o = torch.tensor(origin, dtype=torch.float64, device=dev)
    dz = dirs[:, 2]
    inf = torch.full_like(dz, float("inf"))
    tg = torch.where(dz < -1e-9, -o[2] / dz, inf)
    half = world["extent"] / 2
    hit = o[None, :2] + torch.nan_to_num(tg, posinf=0.0)[:, None] * dirs[:, :2]
    ok = torch.isfinite(tg) & (hit[:, 0].abs() < half) & (hit[:, 1].abs() < half)
    t = torch.where(ok, tg, inf)
    boxes = world["boxes"]
    near = boxes[np.hypot(boxes[:, 0] - origin[0], boxes[:, 1] - origin[1]) < rmax + 30.0]
    inv = 1.0 / dirs
    for bx, by, w, d, h in near:
        lo = torch.tensor([bx - w / 2, by - d / 2, 0.0], dtype=torch.float64, device=dev)
        hi = torch.tensor([bx + w / 2, by + d / 2, h], dtype=torch.float64, device=dev)
        t0 = (lo - o) * inv; t1 = (hi - o) * inv
        tn = torch.minimum(t0, t1).amax(dim=1); tf = torch.maximum(t0, t1).amin(dim=1)
        ok = (tn <= tf) & (tn > 0)
        t = torch.where(ok & (tn < t), tn, t)
    return t


def make_scan_torch(world, seed, n_points, kind="ouster", pos=None, device="cuda", prior_noise=True):
    """make_scan at bench sizes.  -> dict(body: torch (n,3) f32 on `device`, x_true, x_prior, P, kind)."""
    import torch
    rng = np.random.default_rng(0xF4571000 + seed)
    g = torch.Generator(device=device)
    g.manual_seed(0xF4571000 + seed)
    R_true = _rot(rng.uniform(-0.03, 0.03), rng.uniform(-0.03, 0.03), rng.uniform(-np.pi, np.pi))
    t_true = np.array([rng.uniform(-4, 4), rng.uniform(-4, 4), rng.uniform(1.5, 2.2)]) if pos is None else np.asarray(pos, float)
    origin = R_true @ T_LI + t_true
    Rt = torch.tensor(R_true.T, dtype=torch.float64, device=device)
    body = torch.zeros((0, 3), dtype=torch.float64, device=device)
    tries = 0
    while len(body) < n_points:
        m = int((n_points - len(body)) * (1.6 if kind == "avia" else 2.6)) + 256
        if kind == "avia":
            az = torch.deg2rad((torch.rand(m, generator=g, device=device, dtype=torch.float64) * 2 - 1) * 35.2)
            el = torch.deg2rad((torch.rand(m, generator=g, device=device, dtype=torch.float64) * 2 - 1) * 38.6)
            o1 = torch.sort(az, stable=True).indices
            o2 = torch.sort(torch.floor(el[o1] * 40), stable=True).indices
            order = o1[o2]
            az, el = az[order], el[order]
        else:
            cols = m // 128 + 1
            azc = (torch.arange(cols, device=device, dtype=torch.float64) + rng.random()) * (2 * np.pi / cols)
            elr = torch.deg2rad(torch.linspace(-22.5, 22.5, 128, device=device, dtype=torch.float64))
            az = azc.repeat_interleave(128)[:m]; el = elr.repeat(cols)[:m]
            az = az + torch.randn(m, generator=g, device=device, dtype=torch.float64) * 1e-4
            el = el + torch.randn(m, generator=g, device=device, dtype=torch.float64) * 1e-4
        d_b = torch.stack([torch.cos(el) * torch.cos(az), torch.cos(el) * torch.sin(az), torch.sin(el)], dim=1)
        r = _cast_t(origin, d_b @ Rt, world)
        ok = torch.isfinite(r) & (r >= 1.0) & (r <= 120.0)
        rr = r[ok] + torch.randn(int(ok.sum()), generator=g, device=device, dtype=torch.float64) * 0.02
        body = torch.cat([body, d_b[ok] * rr[:, None]])
        tries += 1
        if tries > 20:
            raise RuntimeError("scan generator starved")
    body = body[:n_points].float().contiguous()
    x_true = np.zeros(36)
    x_true[0:9] = R_true.reshape(9); x_true[9:12] = t_true
    x_true[12:21] = np.eye(3).reshape(9); x_true[21:24] = T_LI
    x_true[33:36] = [0.0, 0.0, -9.809]
    x_prior = x_true.copy()
    if prior_noise:
        dth = np.deg2rad(rng.uniform(-0.5, 0.5, 3)); dt = rng.uniform(-0.05, 0.05, 3)
        x_prior[0:9] = (R_true @ exp_so3(dth)).reshape(9)
        x_prior[9:12] = t_true + dt
    return dict(body=body, x_true=x_true, x_prior=x_prior, P=np.eye(23) * 1e-4, kind=kind, seed=seed)


def make_down_scan_torch(world, seed, n_points, leaf, kind="avia", pos=None, device="cuda", prior_noise=True, oversample=48):
    """A scan as the node hands it over: feats_down_body, i.e. voxel-filtered at filter_size_surf (src/laser_mapping.cpp:129-130)
    — at most one point per voxel of the lidar frame.  oversample x n_points raw ray hits (make_scan_torch) are thinned to
    the first hit of every `leaf` voxel; where a pose does not show n_points distinct voxels the leaf shrinks by 15 % until
    it does (the leaf used is returned as s["leaf"]); an evenly spaced subset brings the count to exactly n_points.
    Without this a map built from the first scan holds thousands of points per search cell next to the sensor, which no
    map of the reference pipeline ever does."""
    import torch
    s = make_scan_torch(world, seed, n_points * oversample, kind=kind, pos=pos, device=device, prior_noise=prior_noise)
    b = s["body"]
    for _ in range(40):
        k = torch.floor(b.double() / leaf).long() + (1 << 20)
        key = (k[:, 0] << 42) | (k[:, 1] << 21) | k[:, 2]
        order = torch.sort(key, stable=True).indices
        ks = key[order]
        head = torch.ones(len(ks), dtype=torch.bool, device=b.device)
        head[1:] = ks[1:] != ks[:-1]
        first = torch.sort(order[head]).values          # first hit of every voxel, in the scan's own order
        if len(first) >= n_points:
            pick = first[torch.linspace(0, len(first) - 1, n_points, device=b.device).round().long()]
            s["body"] = b[pick].contiguous()
            s["leaf"] = leaf
            return s
        leaf *= 0.85
    raise RuntimeError("down-sampled scan generator starved")


def make_replay_pool(world, poses, n_points, leaf, kind="avia", device="cuda", seed0=5000, tries=8):
    """One voxel-filtered scan per pose (make_down_scan_torch).  The heading of a synthetic scan is drawn from its seed; a
    heading that faces a nearby wall does not show n_points distinct voxels, so up to `tries` seeds are tried per pose and
    the scan that kept the largest leaf is taken (first one that keeps >= 80 % of the requested leaf)."""
    pool = []
    for i, pos in enumerate(poses):
        best = None
        for t in range(tries):
            s = make_down_scan_torch(world, seed=seed0 + i + 1000 * t, n_points=n_points, leaf=leaf, kind=kind, pos=pos, device=device)
            if best is None or s["leaf"] > best["leaf"]:
                best = s
            if s["leaf"] >= 0.8 * leaf:
                break
        pool.append(best)
    return pool


def street_positions(world, count, pitch=30.0):
    """Sensor positions on street intersections of the box grid (free of buildings), nearest the origin first."""
    nb = max(int(world["extent"] // pitch), 1)
    lines = np.arange(nb + 1) * pitch - nb * pitch / 2
    xs, ys = np.meshgrid(lines, lines)
    p = np.stack([xs.ravel(), ys.ravel()], axis=1)
    lim = world["extent"] / 2 - 40.0
    p = p[(np.abs(p[:, 0]) < lim) & (np.abs(p[:, 1]) < lim)]
    p = p[np.argsort(np.hypot(p[:, 0], p[:, 1]), kind="stable")]
    if len(p) == 0:
        p = np.zeros((1, 2))
    return [np.array([p[i % len(p), 0], p[i % len(p), 1], 1.8]) for i in range(count)]
