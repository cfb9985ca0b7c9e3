"""TEST INFRASTRUCTURE — ctypes bindings for the CPU oracle (oracle/liboracle.so) and, when built,
the reference's own ikd-Tree (oracle/_ref/libflivo_ref.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_PATH = os.path.join(HERE, "_ref", "libflivo_ref.so")
IVOX_REF_PATH = os.path.join(HERE, "_ref", "libivox_ref.so")

f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")

MODE_V3, MODE_V1 = 0, 1


class State(C.Structure):
    """Mirror of orc_state / flivo_state: row-major rotations."""
    _fields_ = [("rot", C.c_double * 9), ("pos", C.c_double * 3), ("offR", C.c_double * 9),
                ("offT", C.c_double * 3), ("vel", C.c_double * 3), ("bg", C.c_double * 3),
                ("ba", C.c_double * 3), ("grav", C.c_double * 3)]

    @staticmethod
    def make(R=None, t=None, offR=None, offT=None, vel=None, bg=None, ba=None, grav=None):
        s = State()
        s.rot[:] = list(np.asarray(np.eye(3) if R is None else R, dtype=np.float64).reshape(9))
        s.pos[:] = list(np.zeros(3) if t is None else np.asarray(t, dtype=np.float64))
        s.offR[:] = list(np.asarray(np.eye(3) if offR is None else offR, dtype=np.float64).reshape(9))
        s.offT[:] = list(np.zeros(3) if offT is None else np.asarray(offT, dtype=np.float64))
        s.vel[:] = list(np.zeros(3) if vel is None else np.asarray(vel, dtype=np.float64))
        s.bg[:] = list(np.zeros(3) if bg is None else np.asarray(bg, dtype=np.float64))
        s.ba[:] = list(np.zeros(3) if ba is None else np.asarray(ba, dtype=np.float64))
        s.grav[:] = list(np.array([0.0, 0.0, -9.809]) if grav is None else np.asarray(grav, dtype=np.float64))
        return s

    def copy(self):
        s = State()
        C.memmove(C.byref(s), C.byref(self), C.sizeof(State))
        return s

    def as_array(self):
        return np.frombuffer(bytes(self), dtype=np.float64).copy()

    @staticmethod
    def from_array(a):
        s = State()
        a = np.ascontiguousarray(a, dtype=np.float64)
        assert a.size == 36
        C.memmove(C.byref(s), a.ctypes.data, C.sizeof(State))
        return s


def build(force=False):
    """Compile the oracle (and oracle/_ref when /root/reference is present). Building the checker
    is not using it."""
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < max(
            os.path.getmtime(os.path.join(HERE, f)) for f in ("flivo_oracle.cpp", "ivox_oracle.cpp", "flivo_oracle.h")):
        subprocess.check_call(["make", "-C", HERE, os.path.join(HERE, "liboracle.so")],
                              stdout=subprocess.DEVNULL)
    ref_src = "/root/reference/include/ikd-Tree/ikd_Tree.cpp"
    if os.path.exists(ref_src):
        newest = max(os.path.getmtime(os.path.join(HERE, f)) for f in
                     ("flivo_oracle.cpp", "ikd_ref_wrap.cpp", "flivo_oracle.h"))
        ivox_newest = max([os.path.getmtime(os.path.join(HERE, "ivox_ref_wrap.cpp"))] +
                          [os.path.getmtime(os.path.join(dp, f)) for dp, _, fs in os.walk(os.path.join(HERE, "shim")) for f in fs])
        if (force or not os.path.exists(REF_PATH) or os.path.getmtime(REF_PATH) < newest
                or not os.path.exists(IVOX_REF_PATH) or os.path.getmtime(IVOX_REF_PATH) < ivox_newest):
            subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        L.orc_body_to_world.argtypes = [f32p, C.c_size_t, C.c_size_t, C.POINTER(State), f32p]
        L.orc_knn_bruteforce.argtypes = [f32p, C.c_void_p, C.c_void_p, C.c_size_t, f32p, C.c_size_t, C.c_int,
                                         C.c_double, i32p, f32p, f32p, i32p, C.POINTER(C.c_int64)]
        L.orc_esti_plane.argtypes = [f32p, C.c_float, f32p]
        L.orc_esti_plane.restype = C.c_int
        L.orc_h_share_model.argtypes = [f32p, C.c_size_t, C.c_size_t, C.POINTER(State), f32p, i32p, f32p,
                                        C.c_int, C.c_int, u8p, u8p, f32p, f32p, f64p, f64p,
                                        C.POINTER(C.c_int64), C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        L.orc_repack_v1.argtypes = [f64p, f64p, C.c_double, f64p, f64p]
        L.orc_solve_v3.argtypes = [C.POINTER(State), C.POINTER(State), f64p, f64p, f64p, C.c_double, C.c_int,
                                   C.c_int, f64p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                   f64p, f64p]
        L.orc_solve_v1.argtypes = [C.POINTER(State), C.POINTER(State), f64p, f64p, f64p, C.c_int, C.c_int,
                                   C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), f64p]
        L.orc_map_create.argtypes = [C.c_float]
        L.orc_map_create.restype = C.c_void_p
        L.orc_map_destroy.argtypes = [C.c_void_p]
        L.orc_map_build.argtypes = [C.c_void_p, f32p, C.c_size_t]
        L.orc_map_add_points.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_int]
        L.orc_map_add_points.restype = C.c_int
        L.orc_map_delete_boxes.argtypes = [C.c_void_p, f32p, C.c_size_t]
        L.orc_map_delete_boxes.restype = C.c_int
        L.orc_map_live.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_map_live.restype = C.c_size_t
        L.orc_undistort.argtypes = [f32p, C.c_size_t, C.c_size_t, f32p, f64p, C.c_int, f64p, f64p, f64p, f64p]
        L.orc_voxel_grid.argtypes = [f32p, C.c_void_p, C.c_size_t, C.c_float, f32p, C.c_void_p, C.c_size_t]
        L.orc_voxel_grid.restype = C.c_long
        L.orc_ivox_create.argtypes = [C.c_float, C.c_int, C.c_size_t]
        L.orc_ivox_create.restype = C.c_void_p
        L.orc_ivox_destroy.argtypes = [C.c_void_p]
        L.orc_ivox_add_points.argtypes = [C.c_void_p, f32p, C.c_size_t]
        L.orc_ivox_num_grids.argtypes = [C.c_void_p]
        L.orc_ivox_num_grids.restype = C.c_size_t
        L.orc_ivox_evicted.argtypes = [C.c_void_p]
        L.orc_ivox_evicted.restype = C.c_size_t
        L.orc_ivox_points.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_ivox_points.restype = C.c_size_t
        L.orc_ivox_knn.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_int, C.c_double, i32p, f32p, f32p, i32p,
                                   C.POINTER(C.c_int64)]
        L.orc_ivox_map_filter.argtypes = [f32p, C.c_size_t, f32p, i32p, C.c_double, u8p]
        L.orc_gnss_block.argtypes = [f64p, f64p, f64p, f64p, f64p, f64p, f64p]
        _lib = L
    return _lib


def have_ref():
    build()
    return os.path.exists(REF_PATH)


def ref():
    global _ref
    if _ref is None:
        build()
        L = C.CDLL(REF_PATH)
        L.ref_tree_create.argtypes = [C.c_float]
        L.ref_tree_create.restype = C.c_void_p
        L.ref_tree_destroy.argtypes = [C.c_void_p]
        L.ref_tree_build.argtypes = [C.c_void_p, f32p, C.c_size_t]
        L.ref_tree_add_points.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_int]
        L.ref_tree_add_points.restype = C.c_int
        L.ref_tree_delete_boxes.argtypes = [C.c_void_p, f32p, C.c_size_t]
        L.ref_tree_delete_boxes.restype = C.c_int
        L.ref_tree_size.argtypes = [C.c_void_p]
        L.ref_tree_validnum.argtypes = [C.c_void_p]
        L.ref_tree_flatten.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.ref_tree_flatten.restype = C.c_size_t
        L.ref_tree_knn.argtypes = [C.c_void_p, f32p, C.c_size_t, C.c_int, C.c_double, C.c_int,
                                   i32p, f32p, C.c_void_p, i32p]
        L.ref_iekf_update.argtypes = [C.c_void_p, f32p, C.c_size_t, C.POINTER(State), f64p, C.c_int, C.c_int,
                                      C.c_double, f64p, C.c_int, C.c_int, f64p, C.POINTER(C.c_int),
                                      C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]
        L.ref_iekf_update.restype = C.c_int
        _ref = L
    return _ref


# ------------------------------------------------------------------------------------------------
# numpy-level helpers
# ------------------------------------------------------------------------------------------------
def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def body_to_world(body, st):
    body = _c(body, np.float32)
    out = np.empty((body.shape[0], 3), np.float32)
    lib().orc_body_to_world(body, body.shape[0], body.shape[1], C.byref(st), out)
    return out


def knn_bruteforce(map_xyz, queries, k=5, max_dist=np.inf, ids=None, alive=None):
    map_xyz = _c(map_xyz, np.float32)
    q = _c(queries, np.float32)
    nq = q.shape[0]
    oid = np.empty((nq, k), np.int32)
    osq = np.empty((nq, k), np.float32)
    oxyz = np.empty((nq, k, 3), np.float32)
    ocnt = np.empty(nq, np.int32)
    ties = C.c_int64(0)
    idp = None if ids is None else _c(ids, np.int32).ctypes.data
    alp = None if alive is None else _c(alive, np.uint8).ctypes.data
    keep = (ids, alive)
    lib().orc_knn_bruteforce(map_xyz, idp, alp, map_xyz.shape[0], q, nq, k, float(max_dist),
                             oid, osq, oxyz, ocnt, C.byref(ties))
    del keep
    return oid, osq, oxyz, ocnt, ties.value


def esti_plane(near5, threshold=0.1):
    out = np.zeros(4, np.float32)
    ok = lib().orc_esti_plane(_c(near5, np.float32).reshape(15), threshold, out)
    return bool(ok), out


def h_share_model(body, st, near_xyz, near_cnt, near_d5, searched, mode=MODE_V3, mask_prev=None):
    body = _c(body, np.float32)
    n = body.shape[0]
    mask = np.ones(n, np.uint8) if mask_prev is None else _c(mask_prev, np.uint8).copy()
    eff = np.zeros(n, np.uint8)
    normvec = np.zeros((n, 4), np.float32)
    world = np.zeros((n, 3), np.float32)
    HTH = np.zeros(144)
    HTz = np.zeros(12)
    n_eff = C.c_int64(0)
    rs = C.c_double(0)
    border = C.c_int64(0)
    lib().orc_h_share_model(body, n, body.shape[1], C.byref(st), _c(near_xyz, np.float32).reshape(-1),
                            _c(near_cnt, np.int32), _c(near_d5, np.float32), int(searched), int(mode),
                            mask, eff, normvec.reshape(-1), world.reshape(-1), HTH, HTz,
                            C.byref(n_eff), C.byref(rs), C.byref(border))
    return dict(mask=mask, eff=eff, normvec=normvec, world=world, HTH=HTH.reshape(12, 12), HTz=HTz,
                n_eff=n_eff.value, res_sum=rs.value, borderline=border.value)


def repack_v1(HTH, HTz, laser_point_cov=0.001):
    HPH = np.zeros(81)
    HPL = np.zeros(9)
    lib().orc_repack_v1(_c(HTH, np.float64).reshape(-1), _c(HTz, np.float64), laser_point_cov, HPH, HPL)
    return HPH.reshape(9, 9), HPL


def solve_v3(x, x_prop, P_prop, HTH, HTz, R, it, max_iter, limit, t):
    tt = C.c_int(t)
    conv = C.c_int(0)
    stop = C.c_int(0)
    dx = np.zeros(23)
    P_out = np.zeros(23 * 23)
    lib().orc_solve_v3(C.byref(x), C.byref(x_prop), _c(P_prop, np.float64).reshape(-1),
                       _c(HTH, np.float64).reshape(-1), _c(HTz, np.float64), R, it, max_iter,
                       _c(limit, np.float64), C.byref(tt), C.byref(conv), C.byref(stop), dx, P_out)
    return dict(t=tt.value, converge=conv.value, stopped=stop.value, dx=dx, P=P_out.reshape(23, 23))


def solve_v1(x, x_prop, cov, HPH, HPL, it, max_iter, rematch):
    rm = C.c_int(rematch)
    sn = C.c_int(0)
    stop = C.c_int(0)
    sol = np.zeros(18)
    cov = _c(cov, np.float64).reshape(-1).copy()
    lib().orc_solve_v1(C.byref(x), C.byref(x_prop), cov, _c(HPH, np.float64).reshape(-1), _c(HPL, np.float64),
                       it, max_iter, C.byref(rm), C.byref(sn), C.byref(stop), sol)
    return dict(rematch=rm.value, search_next=sn.value, stopped=stop.value, solution=sol, cov=cov.reshape(18, 18))


class OracleMap:
    """Restated Add_Points / Delete_Point_Boxes semantics over a flat list."""

    def __init__(self, ds):
        self.h = lib().orc_map_create(ds)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_map_destroy(self.h)
            self.h = None

    def build(self, xyz):
        xyz = _c(xyz, np.float32)
        lib().orc_map_build(self.h, xyz, xyz.shape[0])

    def add_points(self, xyz, downsample_on=True):
        xyz = _c(xyz, np.float32)
        return lib().orc_map_add_points(self.h, xyz, xyz.shape[0], int(downsample_on))

    def delete_boxes(self, boxes):
        boxes = _c(boxes, np.float32).reshape(-1, 6)
        return lib().orc_map_delete_boxes(self.h, boxes, boxes.shape[0])

    def live(self):
        n = lib().orc_map_live(self.h, None, None, 0)
        xyz = np.empty((n, 3), np.float32)
        ids = np.empty(n, np.int32)
        lib().orc_map_live(self.h, xyz.ctypes.data, ids.ctypes.data, n)
        return xyz, ids


class OracleIVox:
    """Restated faster_lio::IVox<3, DEFAULT> (oracle/ivox_oracle.cpp)."""

    def __init__(self, resolution, nearby=19, capacity=0):
        self.h = lib().orc_ivox_create(resolution, nearby, capacity)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_ivox_destroy(self.h)
            self.h = None

    def add_points(self, xyz):
        xyz = _c(xyz, np.float32)
        lib().orc_ivox_add_points(self.h, xyz, xyz.shape[0])

    def num_grids(self):
        return lib().orc_ivox_num_grids(self.h)

    def evicted(self):
        return lib().orc_ivox_evicted(self.h)

    def points(self):
        n = lib().orc_ivox_points(self.h, None, None, 0)
        xyz = np.empty((n, 3), np.float32)
        ids = np.empty(n, np.int32)
        lib().orc_ivox_points(self.h, xyz.ctypes.data, ids.ctypes.data, n)
        return xyz, ids

    def knn(self, queries, k=5, max_range=5.0):
        """-> (ids, sqd, xyz, count, ties); count = -1 where GetClosestPoint found no candidate."""
        q = _c(queries, np.float32)
        n = q.shape[0]
        oid = np.empty((n, k), np.int32)
        osq = np.empty((n, k), np.float32)
        oxyz = np.empty((n, k, 3), np.float32)
        ocnt = np.empty(n, np.int32)
        ties = C.c_int64(0)
        lib().orc_ivox_knn(self.h, q, n, k, float(max_range), oid, osq, oxyz, ocnt, C.byref(ties))
        return oid, osq, oxyz, ocnt, ties.value


def ivox_map_filter(world_xyz, near_xyz, near_cnt, filter_size_map_min):
    w = _c(world_xyz, np.float32)
    flags = np.empty(w.shape[0], np.uint8)
    lib().orc_ivox_map_filter(w, w.shape[0], _c(near_xyz, np.float32), _c(near_cnt, np.int32), float(filter_size_map_min), flags)
    return flags


def gnss_block(st, gnss_pos, antlever, std):
    """Restated GNSSProcessing::computeH -> (HTH 6x6, HTL 6) at state st."""
    x = st.as_array()
    HTH, HTL = np.zeros(36), np.zeros(6)
    lib().orc_gnss_block(_c(x[0:9], np.float64), _c(x[9:12], np.float64), _c(gnss_pos, np.float64), _c(antlever, np.float64),
                         _c(std, np.float64), HTH, HTL)
    return HTH.reshape(6, 6), HTL


_ivox_ref = None


def have_ivox_ref():
    build()
    return os.path.exists(IVOX_REF_PATH)


def ivox_ref():
    global _ivox_ref
    if _ivox_ref is None:
        build()
        L = C.CDLL(IVOX_REF_PATH)
        L.ref_ivox_create.argtypes = [C.c_float, C.c_int, C.c_longlong]
        L.ref_ivox_create.restype = C.c_void_p
        L.ref_ivox_destroy.argtypes = [C.c_void_p]
        L.ref_ivox_add_points.argtypes = [C.c_void_p, f32p, C.c_longlong]
        L.ref_ivox_num_grids.argtypes = [C.c_void_p]
        L.ref_ivox_num_grids.restype = C.c_longlong
        L.ref_ivox_knn.argtypes = [C.c_void_p, f32p, C.c_longlong, C.c_int, C.c_double, f32p, i32p]
        _ivox_ref = L
    return _ivox_ref


class RefIVox:
    """The reference's own faster_lio::IVox<3, DEFAULT> (include/ivox3d headers compiled unmodified into
    oracle/_ref/libivox_ref.so against the Eigen / pcl stand-ins of oracle/shim).  nearby: 1 / 7 / 19 / 27."""

    def __init__(self, resolution, nearby=19, capacity=1000000):
        self.h = ivox_ref().ref_ivox_create(resolution, {1: 0, 7: 1, 19: 2, 27: 3}[nearby], capacity)

    def __del__(self):
        if getattr(self, "h", None):
            ivox_ref().ref_ivox_destroy(self.h)
            self.h = None

    def add_points(self, xyz):
        xyz = _c(xyz, np.float32)
        ivox_ref().ref_ivox_add_points(self.h, xyz, xyz.shape[0])

    def num_grids(self):
        return ivox_ref().ref_ivox_num_grids(self.h)

    def knn(self, queries, k=5, max_range=5.0):
        """-> (xyz [n][k][3] in the order GetClosestPoint returned them, count; -1 = returned false)."""
        q = _c(queries, np.float32)
        n = q.shape[0]
        oxyz = np.zeros((n, k, 3), np.float32)
        ocnt = np.empty(n, np.int32)
        ivox_ref().ref_ivox_knn(self.h, q, n, k, float(max_range), oxyz, ocnt)
        return oxyz, ocnt


class RefTree:
    """The reference's own KD_TREE (compiled unmodified into oracle/_ref)."""

    def __init__(self, ds):
        self.h = ref().ref_tree_create(ds)

    def close(self):
        if getattr(self, "h", None):
            ref().ref_tree_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def build(self, xyz):
        xyz = _c(xyz, np.float32)
        ref().ref_tree_build(self.h, xyz, xyz.shape[0])

    def add_points(self, xyz, downsample_on=True):
        xyz = _c(xyz, np.float32)
        return ref().ref_tree_add_points(self.h, xyz, xyz.shape[0], int(downsample_on))

    def delete_boxes(self, boxes):
        boxes = _c(boxes, np.float32).reshape(-1, 6)
        return ref().ref_tree_delete_boxes(self.h, boxes, boxes.shape[0])

    def size(self):
        return ref().ref_tree_size(self.h)

    def validnum(self):
        return ref().ref_tree_validnum(self.h)

    def flatten(self):
        n = ref().ref_tree_flatten(self.h, None, None, 0)
        xyz = np.empty((n, 3), np.float32)
        ids = np.empty(n, np.int32)
        ref().ref_tree_flatten(self.h, xyz.ctypes.data, ids.ctypes.data, n)
        return xyz, ids

    def knn(self, queries, k=5, max_dist=np.inf, nthreads=0, want_xyz=True):
        q = _c(queries, np.float32)
        nq = q.shape[0]
        oid = np.empty((nq, k), np.int32)
        osq = np.empty((nq, k), np.float32)
        oxyz = np.empty((nq, k, 3), np.float32) if want_xyz else None
        ocnt = np.empty(nq, np.int32)
        ref().ref_tree_knn(self.h, q, nq, k, float(max_dist), nthreads, oid, osq,
                           None if oxyz is None else oxyz.ctypes.data, ocnt)
        return oid, osq, oxyz, ocnt

    def iekf_update(self, body, x, P, mode=MODE_V3, max_iter=4, laser_point_cov=0.001, limit=None,
                    nthreads=0, force_search=False, want_outputs=False):
        body = _c(body, np.float32)
        n = body.shape[0]
        assert body.shape[1] == 3
        limit = np.full(23, 0.001) if limit is None else _c(limit, np.float64)
        P = _c(P, np.float64).copy()
        x = x.copy()
        timings = np.zeros(3)
        searches = C.c_int(0)
        n_eff = C.c_int64(0)
        eff = np.zeros(n, np.uint8) if want_outputs else None
        nid = np.zeros((n, 5), np.int32) if want_outputs else None
        trips = ref().ref_iekf_update(self.h, body, n, C.byref(x), P.reshape(-1), mode, max_iter,
                                      laser_point_cov, limit, nthreads, int(force_search), timings,
                                      C.byref(searches), C.byref(n_eff),
                                      None if eff is None else eff.ctypes.data,
                                      None if nid is None else nid.ctypes.data)
        return dict(x=x, P=P, trips=trips, searches=searches.value, n_eff=n_eff.value, timings=timings,
                    eff=eff, near_id=nid)


# ---- front end ----
def undistort(xyz, t_ms, poses, rot_end, pos_end, R_LI, t_LI):
    """ImuProcess::UndistortPcl backward loop.  xyz (n,3) float32, t_ms (n,) float32 ascending, poses (m,22) float64
    rows [offset_time, acc3, gyr3, vel3, pos3, rot9].  Returns the compensated (n,3) float32."""
    out = np.ascontiguousarray(xyz, np.float32).copy()
    t = _c(t_ms, np.float32)
    P = _c(poses, np.float64)
    a, b, c, d = (_c(np.asarray(v).reshape(-1), np.float64) for v in (rot_end, pos_end, R_LI, t_LI))
    lib().orc_undistort(out, out.shape[0], 3, t, P.reshape(-1), P.shape[0], a, b, c, d)
    return out


def voxel_grid(xyzi, leaf, t_ms=None):
    """pcl::VoxelGrid::filter.  xyzi (n,4) float32.  Returns (out (m,4) float32, out_t (m,) float32 or None)."""
    x = _c(xyzi, np.float32)
    n = x.shape[0]
    out = np.zeros((max(n, 1), 4), np.float32)
    t = _c(t_ms, np.float32) if t_ms is not None else None
    ot = np.zeros(max(n, 1), np.float32) if t is not None else None
    m = lib().orc_voxel_grid(x, t.ctypes.data if t is not None else None, n, float(leaf),
                             out, ot.ctypes.data if ot is not None else None, n)
    return out[:m].copy(), (ot[:m].copy() if ot is not None else None)
