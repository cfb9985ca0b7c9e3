"""fast-livo-noted_b200 — B200-native (sm_100a) LIO measurement-update hot path of FAST-LIVO.

The product is libflivo_b200.so (C ABI: include/flivo.h; CUDA sources: csrc/).  The reference is C++, so the
drop-in host classes are C++ (host/ikd_Tree.h, host/laser_mapping.h); this module is only the ctypes view of
the same C ABI used by tests/ and bench.py.  There is no CPU fallback: if the library is missing or no CUDA
device exists, construction raises.

The directory name carries hyphens (fixed by the task layout), so import it with
`__graft_entry__.load_package()`.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libflivo_b200.so")

MODE_V3, MODE_V1 = 0, 1
K = 5
OK, E_CUDA, E_ARG, E_CAPACITY, E_NCCL, E_STATE = 0, -1, -2, -3, -4, -5


class FlivoError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("flivo error %d: %s" % (code, msg))
        self.code = code


class State(C.Structure):
    """flivo_state (include/flivo.h): row-major rotations; superset of StatesGroup / state_ikfom."""
    _fields_ = [("rot", C.c_double * 9), ("pos", C.c_double * 3), ("offR", C.c_double * 9),
                ("offT", C.c_double * 3), ("vel", C.c_double * 3), ("bg", C.c_double * 3),
                ("ba", C.c_double * 3), ("grav", C.c_double * 3)]

    @staticmethod
    def from_array(a):
        s = State()
        a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1)
        assert a.size == 36
        C.memmove(C.byref(s), a.ctypes.data, C.sizeof(State))
        return s

    @staticmethod
    def make(R=None, t=None, offR=None, offT=None, vel=None, bg=None, ba=None, grav=None):
        a = np.zeros(36)
        a[0:9] = np.asarray(np.eye(3) if R is None else R, dtype=np.float64).reshape(9)
        a[9:12] = 0 if t is None else t
        a[12:21] = np.asarray(np.eye(3) if offR is None else offR, dtype=np.float64).reshape(9)
        a[21:24] = 0 if offT is None else offT
        a[24:27] = 0 if vel is None else vel
        a[27:30] = 0 if bg is None else bg
        a[30:33] = 0 if ba is None else ba
        a[33:36] = [0.0, 0.0, -9.809] if grav is None else grav
        return State.from_array(a)

    def as_array(self):
        return np.frombuffer(bytes(self), dtype=np.float64).copy()

    def copy(self):
        return State.from_array(self.as_array())


IVOX_OFF, IVOX_CENTER, IVOX_NEARBY6, IVOX_NEARBY18, IVOX_NEARBY26 = 0, 1, 7, 19, 27


class Options(C.Structure):
    _fields_ = [("downsample_size", C.c_float), ("cell_ratio", C.c_int32), ("initial_points", C.c_uint64),
                ("sort_scan", C.c_int32), ("ivox_nearby", C.c_int32)]


class Cfg(C.Structure):
    _fields_ = [("mode", C.c_int32), ("max_iter", C.c_int32), ("laser_point_cov", C.c_double),
                ("limit", C.c_double * 23), ("force_search", C.c_int32), ("flags", C.c_int32)]


class Report(C.Structure):
    _fields_ = [("trips", C.c_int32), ("searches", C.c_int32), ("n_eff", C.c_int32), ("converged", C.c_int32),
                ("res_mean", C.c_double), ("dx_last", C.c_double * 23), ("gpu_ms", C.c_float),
                ("kernel_launches", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("map_live", C.c_int64), ("map_cells", C.c_int64), ("map_blocks", C.c_int64),
                ("dir_capacity", C.c_int64), ("bucket_capacity", C.c_int64), ("last_far_queries", C.c_int64),
                ("last_candidates", C.c_int64), ("last_fused_ms", C.c_float * 16), ("last_solve_ms", C.c_float * 16),
                ("last_trip_searched", C.c_int32 * 16)]


class Pose6D(C.Structure):
    """flivo_pose6d (msg/Pose6D.msg)."""
    _fields_ = [("offset_time", C.c_double), ("acc", C.c_double * 3), ("gyr", C.c_double * 3), ("vel", C.c_double * 3),
                ("pos", C.c_double * 3), ("rot", C.c_double * 9)]


# every symbol include/flivo.h declares (tests check the library exports all of them)
SYMBOLS = [
    "flivo_ctx_create", "flivo_ctx_destroy", "flivo_last_error", "flivo_version", "flivo_ctx_stream", "flivo_sync",
    "flivo_map_set_downsample", "flivo_map_build", "flivo_map_add_points", "flivo_map_delete_boxes",
    "flivo_map_add_boxes", "flivo_map_delete_points", "flivo_map_size", "flivo_map_range", "flivo_map_flatten",
    "flivo_map_acquire_removed", "flivo_knn", "flivo_scan_upload", "flivo_scan_set_device", "flivo_h_share_model",
    "flivo_repack_v1", "flivo_iekf_update", "flivo_map_add_scan", "flivo_comm_unique_id", "flivo_comm_init",
    "flivo_comm_destroy", "flivo_get_stats", "flivo_debug_solve_stamps", "flivo_debug_search_stamps", "flivo_debug_scan_order",
    "flivo_raw_upload", "flivo_raw_set_device", "flivo_raw_undistort", "flivo_raw_downsample", "flivo_raw_download",
    "flivo_scan_download", "flivo_peer_export", "flivo_peer_attach", "flivo_peer_detach", "flivo_set_gnss",
    "flivo_peer_local_base", "flivo_peer_attach_local", "flivo_peer_set_timeout", "flivo_peer_reset",
    "flivo_scan_upload_async", "flivo_scan_commit", "flivo_host_register", "flivo_host_unregister",
    "flivo_iekf_update_begin", "flivo_iekf_update_end", "flivo_knn_noalloc", "flivo_frame",
]

_lib = None


def lib():
    """Load libflivo_b200.so (building it first when a toolkit is present).  Raises if it cannot be had."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.build()
    alt = os.environ.get("FLIVO_LIB")   # development only: A/B another build of the same ABI on one GPU box
    if alt:
        path = alt if os.path.isabs(alt) else os.path.join(HERE, alt)
    L = C.CDLL(path)
    vp, sz, i32, f32 = C.c_void_p, C.c_size_t, C.c_int32, C.c_float
    L.flivo_ctx_create.argtypes = [C.c_int, C.POINTER(Options), C.POINTER(vp)]
    L.flivo_ctx_destroy.argtypes = [vp]
    L.flivo_ctx_destroy.restype = None
    L.flivo_last_error.argtypes = [vp]
    L.flivo_last_error.restype = C.c_char_p
    L.flivo_version.restype = C.c_char_p
    L.flivo_ctx_stream.argtypes = [vp]
    L.flivo_ctx_stream.restype = vp
    L.flivo_sync.argtypes = [vp]
    L.flivo_map_set_downsample.argtypes = [vp, f32]
    L.flivo_map_build.argtypes = [vp, vp, sz, sz]
    L.flivo_map_add_points.argtypes = [vp, vp, sz, sz, C.c_int, C.POINTER(i32)]
    L.flivo_map_delete_boxes.argtypes = [vp, vp, sz, C.POINTER(i32)]
    L.flivo_map_add_boxes.argtypes = [vp, vp, sz, C.POINTER(i32)]
    L.flivo_map_delete_points.argtypes = [vp, vp, sz, sz, C.POINTER(i32)]
    L.flivo_map_size.argtypes = [vp, C.POINTER(C.c_int64)]
    L.flivo_map_range.argtypes = [vp, vp]
    L.flivo_map_flatten.argtypes = [vp, vp, vp, sz, C.POINTER(sz)]
    L.flivo_map_acquire_removed.argtypes = [vp, vp, sz, C.POINTER(sz)]
    L.flivo_knn.argtypes = [vp, vp, sz, sz, C.c_int, C.c_double, vp, vp, vp, vp]
    L.flivo_knn_noalloc.argtypes = [vp, vp, sz, sz, C.c_int, C.c_double, vp, vp, vp, vp]
    L.flivo_scan_upload.argtypes = [vp, vp, sz, sz]
    L.flivo_scan_set_device.argtypes = [vp, vp, sz]
    L.flivo_scan_upload_async.argtypes = [vp, vp, sz, sz]
    L.flivo_scan_commit.argtypes = [vp]
    L.flivo_host_register.argtypes = [vp, sz]
    L.flivo_host_unregister.argtypes = [vp]
    L.flivo_iekf_update_begin.argtypes = [vp, C.POINTER(State), vp, C.POINTER(Cfg)]
    L.flivo_iekf_update_end.argtypes = [vp, C.POINTER(State), vp, C.POINTER(Report)]
    L.flivo_frame.argtypes = [vp, C.POINTER(State), vp, C.POINTER(Cfg), vp, sz, sz, C.POINTER(Report), C.POINTER(i32)]
    L.flivo_h_share_model.argtypes = [vp, C.POINTER(State), C.c_int, C.c_int, vp, vp, C.POINTER(i32),
                                      C.POINTER(C.c_double), vp, vp, vp, vp, vp, vp]
    L.flivo_repack_v1.argtypes = [vp, vp, C.c_double, vp, vp]
    L.flivo_iekf_update.argtypes = [vp, C.POINTER(State), vp, C.POINTER(Cfg), C.POINTER(Report)]
    L.flivo_map_add_scan.argtypes = [vp, C.POINTER(State), C.POINTER(i32)]
    L.flivo_set_gnss.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.flivo_comm_unique_id.argtypes = [vp]
    L.flivo_comm_init.argtypes = [vp, vp, C.c_int, C.c_int]
    L.flivo_comm_destroy.argtypes = [vp]
    L.flivo_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.flivo_debug_solve_stamps.argtypes = [vp, C.POINTER(C.c_uint64), C.c_int]
    L.flivo_debug_search_stamps.argtypes = [vp, C.POINTER(C.c_uint64), C.c_int, C.POINTER(C.c_int)]
    L.flivo_debug_scan_order.argtypes = [vp, C.POINTER(C.c_uint32), C.c_size_t]
    L.flivo_peer_export.argtypes = [vp, vp]
    L.flivo_peer_attach.argtypes = [vp, vp, C.c_int, C.c_int]
    L.flivo_peer_detach.argtypes = [vp]
    L.flivo_peer_local_base.argtypes = [vp, C.POINTER(vp)]
    L.flivo_peer_attach_local.argtypes = [vp, C.POINTER(vp), C.c_int, C.c_int]
    L.flivo_peer_set_timeout.argtypes = [vp, C.c_double]
    L.flivo_peer_reset.argtypes = [vp]
    L.flivo_raw_upload.argtypes = [vp, vp, sz, sz, C.c_int, C.c_int]
    L.flivo_raw_set_device.argtypes = [vp, vp, vp, sz]
    L.flivo_raw_undistort.argtypes = [vp, C.POINTER(Pose6D), C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                      C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.flivo_raw_downsample.argtypes = [vp, C.c_float, C.POINTER(sz)]
    L.flivo_raw_download.argtypes = [vp, C.POINTER(C.c_float), sz, C.POINTER(sz)]
    L.flivo_scan_download.argtypes = [vp, C.POINTER(C.c_float), sz, C.POINTER(sz)]
    _lib = L
    return L


def _pts(a):
    """(n,3) or (n,4) float32 C-contiguous array -> (array, n, stride_bytes)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    assert a.ndim == 2 and a.shape[1] in (3, 4), a.shape
    return a, a.shape[0], a.shape[1] * 4


def _ptr(a):
    return None if a is None else a.ctypes.data


CFG_PROFILE = 1
V1_RANGE_CELLS = 24   # FLIVO_V1_RANGE_CELLS


def v1_search_bound_sq(ds, cell_ratio=2):
    """Squared range up to which V1 keeps neighbours (float32 arithmetic of api.cu: search_bound_bits)."""
    edge = np.float32(ds) * np.float32(cell_ratio)
    r = np.float32(V1_RANGE_CELLS) * edge
    return float(max(np.float32(5.0), np.float32(r * r)))


def make_cfg(mode=MODE_V3, max_iter=4, laser_point_cov=0.001, limit=None, force_search=False, profile=False):
    c = Cfg()
    c.mode, c.max_iter, c.laser_point_cov, c.force_search = mode, max_iter, laser_point_cov, int(force_search)
    c.flags = CFG_PROFILE if profile else 0
    lim = np.full(23, 0.001) if limit is None else np.asarray(limit, dtype=np.float64)
    c.limit[:] = list(lim)
    return c


class Flivo:
    """One flivo_ctx (one device + one stream).  Method names follow the C ABI / the reference members."""

    def __init__(self, device=0, downsample_size=0.2, cell_ratio=2, initial_points=0, sort_scan=False, ivox_nearby=0):
        """ivox_nearby = 0: ikd-Tree semantics.  1 / 7 / 19 / 27 (IVOX_*): the map behaves as the reference's
        faster_lio::IVox (include/ivox3d/ivox3d.h) with downsample_size as its voxel resolution."""
        self._L = lib()
        opt = Options(downsample_size, cell_ratio, initial_points, int(sort_scan), int(ivox_nearby))
        h = C.c_void_p()
        rc = self._L.flivo_ctx_create(device, C.byref(opt), C.byref(h))
        if rc != OK:
            raise FlivoError(rc, "flivo_ctx_create failed (no CUDA device? this library has no CPU fallback)")
        self.h = h
        self.n_scan = 0
        self.ds, self.cell_ratio = float(downsample_size), int(cell_ratio) or 2

    def close(self):
        if getattr(self, "h", None):
            self._L.flivo_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc != OK:
            raise FlivoError(rc, (self._L.flivo_last_error(self.h) or b"").decode())

    # ---- map (KD_TREE surface) ----
    def set_downsample_param(self, ds):
        self._ck(self._L.flivo_map_set_downsample(self.h, ds))

    def Build(self, xyz):
        a, n, st = _pts(xyz)
        self._ck(self._L.flivo_map_build(self.h, _ptr(a), n, st))

    def Add_Points(self, xyz, downsample_on=True):
        a, n, st = _pts(xyz)
        out = C.c_int32(0)
        self._ck(self._L.flivo_map_add_points(self.h, _ptr(a), n, st, int(downsample_on), C.byref(out)))
        return out.value

    def Delete_Point_Boxes(self, boxes):
        b = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 6)
        out = C.c_int32(0)
        self._ck(self._L.flivo_map_delete_boxes(self.h, _ptr(b), b.shape[0], C.byref(out)))
        return out.value

    def Add_Point_Boxes(self, boxes):
        b = np.ascontiguousarray(boxes, dtype=np.float32).reshape(-1, 6)
        out = C.c_int32(0)
        self._ck(self._L.flivo_map_add_boxes(self.h, _ptr(b), b.shape[0], C.byref(out)))
        return out.value

    def Delete_Points(self, xyz):
        a, n, st = _pts(xyz)
        out = C.c_int32(0)
        self._ck(self._L.flivo_map_delete_points(self.h, _ptr(a), n, st, C.byref(out)))
        return out.value

    def size(self):
        out = C.c_int64(0)
        self._ck(self._L.flivo_map_size(self.h, C.byref(out)))
        return out.value

    validnum = size

    def tree_range(self):
        out = np.zeros(6, np.float32)
        self._ck(self._L.flivo_map_range(self.h, _ptr(out)))
        return out

    def flatten(self):
        n = C.c_size_t(0)
        self._ck(self._L.flivo_map_flatten(self.h, None, None, 0, C.byref(n)))
        xyz = np.zeros((n.value, 3), np.float32)
        ids = np.zeros(n.value, np.int32)
        if n.value:
            self._ck(self._L.flivo_map_flatten(self.h, _ptr(xyz), _ptr(ids), n.value, C.byref(n)))
        return xyz, ids

    def acquire_removed_points(self):
        n = C.c_size_t(0)
        self._ck(self._L.flivo_map_acquire_removed(self.h, None, 0, C.byref(n)))
        xyz = np.zeros((n.value, 3), np.float32)
        if n.value:
            self._ck(self._L.flivo_map_acquire_removed(self.h, _ptr(xyz), n.value, C.byref(n)))
        return xyz

    def AddPoints(self, xyz):
        """IVox::AddPoints (include/ivox3d/ivox3d.h:247-276): every point is kept."""
        return self.Add_Points(xyz, downsample_on=False)

    def GetClosestPoint(self, queries, max_num=K, max_range=5.0):
        """Batched IVox::GetClosestPoint (include/ivox3d/ivox3d.h:132-204) -> (ids, sqd, xyz, count)."""
        return self.Nearest_Search(queries, max_num, float(max_range) * float(max_range))

    def Nearest_Search(self, queries, k=K, max_dist=np.inf):
        """Batched KD_TREE::Nearest_Search -> (ids, sqd, xyz, count)."""
        a, n, st = _pts(queries)
        oid = np.full((n, k), -1, np.int32)
        osq = np.full((n, k), np.inf, np.float32)
        oxyz = np.zeros((n, k, 3), np.float32)
        ocnt = np.zeros(n, np.int32)
        self._ck(self._L.flivo_knn(self.h, _ptr(a), n, st, k, float(max_dist), _ptr(oxyz), _ptr(osq), _ptr(oid), _ptr(ocnt)))
        return oid, osq, oxyz, ocnt

    def Nearest_Search_noalloc(self, queries, k=K, max_dist=np.inf):
        """flivo_knn_noalloc: the same search through the library's static device scratch (no allocation in the call)."""
        a, n, st = _pts(queries)
        oid = np.full((n, k), -1, np.int32)
        osq = np.full((n, k), np.inf, np.float32)
        oxyz = np.zeros((n, k, 3), np.float32)
        ocnt = np.zeros(n, np.int32)
        self._ck(self._L.flivo_knn_noalloc(self.h, _ptr(a), n, st, k, float(max_dist), _ptr(oxyz), _ptr(osq), _ptr(oid), _ptr(ocnt)))
        return oid, osq, oxyz, ocnt

    # ---- scan + update ----
    def scan_upload(self, body_xyz):
        a, n, st = _pts(body_xyz)
        self._ck(self._L.flivo_scan_upload(self.h, _ptr(a), n, st))
        self.n_scan = n

    def scan_upload_async(self, rec, stride_bytes=None):
        """Stage the NEXT scan (any (n, k) float32 records with xyz first; pinned / registered arrays travel by DMA as they
        are).  The array must stay alive until scan_commit()."""
        assert rec.dtype == np.float32 and rec.flags["C_CONTIGUOUS"] and rec.ndim == 2
        self._next_scan = rec
        self._ck(self._L.flivo_scan_upload_async(self.h, rec.ctypes.data, rec.shape[0], stride_bytes or rec.strides[0]))

    def scan_commit(self):
        self._ck(self._L.flivo_scan_commit(self.h))
        self.n_scan = self._next_scan.shape[0]

    def iekf_update_begin(self, st, P, cfg):
        np_ = 18 if cfg.mode == MODE_V1 else 23
        self._upd = (st.copy(), np.ascontiguousarray(P, dtype=np.float64).reshape(np_, np_).copy(), cfg)
        self._ck(self._L.flivo_iekf_update_begin(self.h, C.byref(self._upd[0]), _ptr(self._upd[1]), C.byref(cfg)))

    def iekf_update_end(self):
        x, Pm, _ = self._upd
        rep = Report()
        self._ck(self._L.flivo_iekf_update_end(self.h, C.byref(x), _ptr(Pm), C.byref(rep)))
        return x, Pm, rep

    def frame(self, st, P, cfg, next_body=None):
        """flivo_frame: commit the staged scan, update, stage `next_body` meanwhile, map_incremental.
        Returns (state, P, Report, points added)."""
        x = st.copy()
        np_ = 18 if cfg.mode == MODE_V1 else 23
        Pm = np.ascontiguousarray(P, dtype=np.float64).reshape(np_, np_).copy()
        rep, added = Report(), C.c_int32(0)
        n_cur = self._next_scan.shape[0]
        if next_body is not None:
            assert next_body.dtype == np.float32 and next_body.flags["C_CONTIGUOUS"] and next_body.ndim == 2
        self._ck(self._L.flivo_frame(self.h, C.byref(x), _ptr(Pm), C.byref(cfg), _ptr(next_body),
                                     0 if next_body is None else next_body.shape[0], 0 if next_body is None else next_body.strides[0],
                                     C.byref(rep), C.byref(added)))
        self.n_scan = n_cur
        if next_body is not None:
            self._next_scan = next_body
        return x, Pm, rep, added.value

    def scan_set_device(self, dev_ptr, n):
        self._ck(self._L.flivo_scan_set_device(self.h, C.c_void_p(dev_ptr), n))
        self.n_scan = n

    def h_share_model(self, st, mode=MODE_V3, do_search=True, want=()):
        """One h_share_model trip.  want: subset of {'selected','eff','nbr_id','nbr_sqd','plane','world'}."""
        n = self.n_scan
        HTH = np.zeros(144)
        HTz = np.zeros(12)
        ne = C.c_int32(0)
        rs = C.c_double(0)
        bufs = dict(selected=np.zeros(n, np.uint8), eff=np.zeros(n, np.uint8), nbr_id=np.zeros((n, 5), np.int32),
                    nbr_sqd=np.zeros((n, 5), np.float32), plane=np.zeros((n, 4), np.float32),
                    world=np.zeros((n, 3), np.float32))
        p = {k: (_ptr(v) if k in want else None) for k, v in bufs.items()}
        self._ck(self._L.flivo_h_share_model(self.h, C.byref(st), mode, int(do_search), _ptr(HTH), _ptr(HTz),
                                             C.byref(ne), C.byref(rs), p["selected"], p["eff"], p["nbr_id"],
                                             p["nbr_sqd"], p["plane"], p["world"]))
        out = dict(HTH=HTH.reshape(12, 12), HTz=HTz, n_eff=ne.value, res_sum=rs.value)
        for k in want:
            out[k] = bufs[k]
        return out

    def repack_v1(self, HTH, HTz, laser_point_cov=0.001):
        HTH = np.ascontiguousarray(HTH, dtype=np.float64).reshape(-1)
        HTz = np.ascontiguousarray(HTz, dtype=np.float64)
        HPH = np.zeros(81)
        HPL = np.zeros(9)
        rc = self._L.flivo_repack_v1(_ptr(HTH), _ptr(HTz), laser_point_cov, _ptr(HPH), _ptr(HPL))
        if rc != OK:
            raise FlivoError(rc, "flivo_repack_v1")
        return HPH.reshape(9, 9), HPL

    def iekf_update(self, st, P, cfg=None, **kw):
        """Whole iterated update on the device.  Returns (state, P, Report)."""
        cfg = cfg or make_cfg(**kw)
        x = st.copy()
        np_ = 18 if cfg.mode == MODE_V1 else 23
        Pm = np.ascontiguousarray(P, dtype=np.float64).reshape(np_, np_).copy()
        rep = Report()
        self._ck(self._L.flivo_iekf_update(self.h, C.byref(x), _ptr(Pm), C.byref(cfg), C.byref(rep)))
        return x, Pm, rep

    def set_gnss(self, pos_global=None, antlever=(0.0, 0.0, 0.0), std=(1.0, 1.0, 1.0)):
        """GNSS observation folded into the following V1 updates (src/laser_mapping.cpp:612-617); None clears it."""
        if pos_global is None:
            self._ck(self._L.flivo_set_gnss(self.h, None, None, None))
            return
        a = (C.c_double * 3)(*[float(v) for v in pos_global])
        b = (C.c_double * 3)(*[float(v) for v in antlever])
        c = (C.c_double * 3)(*[float(v) for v in std])
        self._ck(self._L.flivo_set_gnss(self.h, a, b, c))

    def map_add_scan(self, st):
        out = C.c_int32(0)
        self._ck(self._L.flivo_map_add_scan(self.h, C.byref(st), C.byref(out)))
        return out.value

    def stats(self):
        s = Stats()
        self._ck(self._L.flivo_get_stats(self.h, C.byref(s)))
        return s

    # ---- front end: raw scan -> undistort -> voxel grid (device resident) ----
    def raw_upload(self, rec, time_col=None, intensity_col=None):
        """rec: (n, k) float32 records, xyz in columns 0..2; time_col / intensity_col: column indices or None."""
        rec = np.ascontiguousarray(rec, np.float32)
        self._keep_raw = rec
        self._ck(self._L.flivo_raw_upload(self.h, rec.ctypes.data, rec.shape[0], rec.strides[0],
                                          -1 if time_col is None else 4 * time_col, -1 if intensity_col is None else 4 * intensity_col))

    def raw_undistort(self, poses, rot_end, pos_end, R_LI, t_LI):
        """poses: (m, 22) float64 rows [offset_time, acc3, gyr3, vel3, pos3, rot9]."""
        poses = np.ascontiguousarray(poses, np.float64)
        arr = (Pose6D * poses.shape[0]).from_buffer_copy(poses.tobytes())
        d = lambda a: np.ascontiguousarray(a, np.float64).ctypes.data_as(C.POINTER(C.c_double))
        a, b, c, e = (np.ascontiguousarray(v, np.float64) for v in (rot_end, pos_end, R_LI, t_LI))
        self._ck(self._L.flivo_raw_undistort(self.h, arr, poses.shape[0], d(a), d(b), d(c), d(e)))

    def raw_downsample(self, leaf):
        n = C.c_size_t(0)
        self._ck(self._L.flivo_raw_downsample(self.h, float(leaf), C.byref(n)))
        self.n_scan = n.value
        return n.value

    def _download(self, fn):
        n = C.c_size_t(0)
        self._ck(fn(self.h, None, 0, C.byref(n)))
        out = np.zeros((n.value, 4), np.float32)
        if n.value:
            self._ck(fn(self.h, out.ctypes.data_as(C.POINTER(C.c_float)), n.value, C.byref(n)))
        return out

    def raw_download(self):
        return self._download(self._L.flivo_raw_download)

    def scan_download(self):
        return self._download(self._L.flivo_scan_download)

    def scan_order(self, n):
        """Processing order of the current scan of n points: out[q] = caller index of the point processed q-th."""
        out = np.empty(n, np.uint32)
        self._ck(self._L.flivo_debug_scan_order(self.h, out.ctypes.data_as(C.POINTER(C.c_uint32)), n))
        return out

    def solve_stamps(self, trips=16):
        """[trips][16] device globaltimer stamps (ns) of the solve kernel's phases (profile mode only)."""
        buf = (C.c_uint64 * (trips * 16))()
        self._ck(self._L.flivo_debug_solve_stamps(self.h, buf, trips))
        return np.ctypeslib.as_array(buf).reshape(trips, 16).copy()

    def search_stamps(self):
        """[ctas][4] per-CTA stamps (ns) of the first search trip of the last profile-mode update."""
        cap = 4096
        buf = (C.c_uint64 * ((cap + 8) * 4))()
        n = C.c_int(0)
        self._ck(self._L.flivo_debug_search_stamps(self.h, buf, cap + 8, C.byref(n)))
        a = np.ctypeslib.as_array(buf)
        self.far_totals = a[cap * 4:cap * 4 + 32].copy()   # [sum ns, queries, sum steps, radius histogram 3.., seed-count histogram 16..]
        return a[:cap * 4].reshape(cap, 4)[:min(n.value, cap)].copy()

    def sync(self):
        self._ck(self._L.flivo_sync(self.h))

    def stream(self):
        return self._L.flivo_ctx_stream(self.h)

    # ---- multi-GPU ----
    @staticmethod
    def comm_unique_id():
        buf = (C.c_char * 128)()
        rc = lib().flivo_comm_unique_id(buf)
        if rc != OK:
            raise FlivoError(rc, "flivo_comm_unique_id")
        return bytes(buf)

    def comm_init(self, uid, rank, nranks):
        buf = (C.c_char * 128).from_buffer_copy(uid)
        self._ck(self._L.flivo_comm_init(self.h, buf, rank, nranks))

    def peer_export(self):
        buf = (C.c_char * 64)()
        self._ck(self._L.flivo_peer_export(self.h, buf))
        return bytes(buf)

    def peer_attach(self, handles, rank, nranks):
        """handles: the 64-byte handles of all ranks, in rank order."""
        blob = b"".join(handles)
        buf = (C.c_char * len(blob)).from_buffer_copy(blob)
        self._ck(self._L.flivo_peer_attach(self.h, buf, rank, nranks))

    def peer_detach(self):
        self._ck(self._L.flivo_peer_detach(self.h))

    def peer_local_base(self):
        """Device pointer of this context's exchange buffer (ranks that live in one process)."""
        out = C.c_void_p()
        self._ck(self._L.flivo_peer_local_base(self.h, C.byref(out)))
        return out.value

    def peer_attach_local(self, bases, rank, nranks):
        arr = (C.c_void_p * nranks)(*bases)
        self._ck(self._L.flivo_peer_attach_local(self.h, arr, rank, nranks))

    def peer_set_timeout(self, seconds):
        self._ck(self._L.flivo_peer_set_timeout(self.h, float(seconds)))

    def peer_reset(self):
        self._ck(self._L.flivo_peer_reset(self.h))

    def comm_destroy(self):
        self._ck(self._L.flivo_comm_destroy(self.h))
