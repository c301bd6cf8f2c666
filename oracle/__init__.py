"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE ONLY).

May be imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package (uavmvs_b200)
never imports this module.

  liboracle.so        oracle/uavmvs_oracle.cpp: our restatement (kind "port")
  _ref/libref_acc.so  the reference's libs/acc headers compiled as they lie
                      (kind "reference"; tree build / ray traversal / KD search)
  _ref/libref_cuda_{exact,fast}.so
                      the reference's own kernels (libs/eval/kernels.cu + libs/cacc)
                      compiled for sm_100a as they lie; GPU box only (refcuda())
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

NAI = 0xFFFFFFFF
VIEW_FLOATS = 28  # pos[3], w2c[16], calib[9]

_vp = C.c_void_p
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def build(force=False):
    """Compile liboracle.so (and _ref/ when /root/reference is present)."""
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "uavmvs_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    ref_so = os.path.join(_HERE, "_ref", "libref_acc.so")
    shim = os.path.join(_HERE, "ref_shim", "ref_acc.cpp")
    if os.path.isdir("/root/reference/libs/acc") and (force or not os.path.exists(ref_so)
                                                      or os.path.getmtime(ref_so) < os.path.getmtime(shim)):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
    geom_so = os.path.join(_HERE, "_ref", "libref_geom.so")
    geom_shim = os.path.join(_HERE, "ref_shim", "ref_geom.cpp")
    if os.path.isdir("/root/reference/libs/geom") and (force or not os.path.exists(geom_so)
                                                        or os.path.getmtime(geom_so) < os.path.getmtime(geom_shim)):
        subprocess.check_call(["make", "-C", _HERE, "refgeom"], stdout=subprocess.DEVNULL)
    demo = os.path.join(_HERE, "_ref", "acc_tree_facade")
    demo_src = os.path.join(_HERE, "..", "tests", "cpp", "acc_tree_facade.cpp")
    facade = os.path.join(_HERE, "..", "include", "uavmvs_b200", "cacc.hpp")
    product = os.path.join(_HERE, "..", "uavmvs_b200", "libuavmvs_b200.so")
    if os.path.isdir("/root/reference/libs/acc") and os.path.exists(product) and os.path.exists(demo_src) and (
            force or not os.path.exists(demo) or os.path.getmtime(demo) < max(os.path.getmtime(demo_src), os.path.getmtime(facade))):
        subprocess.check_call(["make", "-C", _HERE, "accdemo"], stdout=subprocess.DEVNULL)
    cu_so = os.path.join(_HERE, "_ref", "libref_cuda_fast.so")
    cu_shim = os.path.join(_HERE, "ref_shim", "ref_cuda.cu")
    if os.path.isdir("/root/reference/libs/eval") and (force or not os.path.exists(cu_so)
                                                       or os.path.getmtime(cu_so) < os.path.getmtime(cu_shim)):
        subprocess.check_call(["make", "-C", _HERE, "refcuda"], stdout=subprocess.DEVNULL)


def _opt(ptr_type):
    """ndpointer that also accepts None."""
    base = ptr_type

    class _Opt(base):
        @classmethod
        def from_param(cls, obj):
            if obj is None:
                return None
            return base.from_param(obj)
    return _Opt


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    build()
    L = C.CDLL(os.path.join(_HERE, "liboracle.so"))
    u32, u64, i32, f32 = C.c_uint32, C.c_uint64, C.c_int, C.c_float
    L.orc_set_threads.restype = i32
    L.orc_set_threads.argtypes = [i32]
    L.orc_bvh_build.restype = _vp
    L.orc_bvh_build.argtypes = [_f32p, u32, _u32p, u32, i32]
    L.orc_bvh_from_arrays.restype = _vp
    L.orc_bvh_from_arrays.argtypes = [_u32p, _f32p, _f32p, _u32p, u32, u32]
    L.orc_bvh_num_nodes.restype = u32
    L.orc_bvh_num_nodes.argtypes = [_vp]
    L.orc_bvh_num_tris.restype = u32
    L.orc_bvh_num_tris.argtypes = [_vp]
    L.orc_bvh_export.argtypes = [_vp, _u32p, _f32p, _f32p, _u32p]
    L.orc_bvh_destroy.argtypes = [_vp]
    L.orc_bvh_intersect.argtypes = [_vp, _f32p, u64, i32, _opt(_u8p), _opt(_u32p), _opt(_f32p), _opt(_u64p)]
    L.orc_kd_build.restype = _vp
    L.orc_bvh_closest_point.argtypes = [_vp, _f32p, u64, f32, _u8p, _u32p, _f32p]
    L.orc_kd_build.argtypes = [_f32p, u32]
    L.orc_kd_from_arrays.restype = _vp
    L.orc_kd_from_arrays.argtypes = [_u32p, _f32p, u32, u32]
    L.orc_kd_num_nodes.restype = u32
    L.orc_kd_num_nodes.argtypes = [_vp]
    L.orc_kd_export.argtypes = [_vp, _u32p, _f32p]
    L.orc_kd_find_nn.argtypes = [_vp, _f32p, u64, _u32p, _opt(_f32p)]
    L.orc_kd_destroy.argtypes = [_vp]
    L.orc_sphere_vertices.restype = u32
    L.orc_sphere_vertices.argtypes = [f32, u32, _opt(_f32p), u32]
    L.orc_update_observation_rays.argtypes = [
        _vp, i32, _f32p, u32, f32, i32, i32, _f32p, _f32p, u32,
        _f32p, _u32p, u32, u64, _opt(_u8p), _opt(_u64p)]
    L.orc_process_observation_rays.argtypes = [_f32p, _u32p, u32, u64, u32]
    L.orc_evaluate_observation_rays.argtypes = [_f32p, _u32p, u32, u64, u32, _f32p, _f32p]
    L.orc_calculate_func_recons.argtypes = [_f32p, f32, _f32p, u32]
    L.orc_reduce.argtypes = [_f32p, u32, _f64p]
    L.orc_populate_spherical_histogram.argtypes = [
        _vp, _vp, _f32p, f32, _f32p, _f32p, _f32p, _f32p, u32, _f64p, _opt(_u32p), _opt(_f32p)]
    L.orc_populate_spherical_histogram_recon.argtypes = [
        _vp, _vp, _f32p, f32, f32, _f32p, _f32p, u32, _f32p, _u32p, u32, u64,
        _f32p, _f32p, _f64p, _opt(_u32p), _opt(_f32p)]
    L.orc_evaluate_spherical_histogram.argtypes = [_f32p, i32, i32, _f32p, u32, _f32p, _f32p, u32, u32]
    L.orc_estimate_capture_difficulty.argtypes = [_vp, f32, u32, _f32p, u32, _f32p, _f32p, u32, _f32p, _f32p]
    L.orc_sample_proxy_cloud.restype = u32
    L.orc_sample_proxy_cloud.argtypes = [_f32p, _u32p, u32, f32, _opt(_f32p), _opt(_f32p), u32]
    _LIB = L
    return L


_REFGEOM = None
REF_OBJECTIVE = C.CFUNCTYPE(C.c_float, C.POINTER(C.c_float), C.c_void_p)


def ref_nelder_mead_step(simplex, objective):
    """libs/opti/nelder_mead.h:29-121 (the reference's own code, oracle/_ref/libref_geom.so) on one Simplex<3>;
    objective(pos[3]) -> float.  Returns (new simplex [4,3], vertex id, value)."""
    R = refgeom()
    s = np.ascontiguousarray(simplex, np.float32).reshape(12).copy()
    cb = REF_OBJECTIVE(lambda p, _u: float(objective(np.array([p[0], p[1], p[2]], np.float32))))
    vid, val = C.c_uint32(0), C.c_float(0)
    R.ref_nelder_mead_step(s, cb, None, C.byref(vid), C.byref(val))
    return s.reshape(4, 3), int(vid.value), float(val.value)



def refgeom():
    """The reference's own libs/geom + libs/util writers (None when oracle/_ref/libref_geom.so was not built)."""
    global _REFGEOM
    if _REFGEOM is not None:
        return _REFGEOM or None
    path = os.path.join(_HERE, "_ref", "libref_geom.so")
    if not os.path.exists(path):
        _REFGEOM = False
        return None
    R = C.CDLL(path)
    u32, i32, f32 = C.c_uint32, C.c_int, C.c_float
    R.ref_sphere_vertices.restype = u32
    R.ref_sphere_vertices.argtypes = [f32, u32, _opt(_f32p), u32]
    R.ref_save_numpy_f32.argtypes = [C.c_char_p, _f32p, u32, u32]
    R.ref_save_numpy_i32.argtypes = [C.c_char_p, np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS"), u32, u32]
    R.ref_save_volume.argtypes = [C.c_char_p, u32, u32, u32, _f32p, _f32p, u32, _u32p, i32, i32, i32, _f32p]
    R.ref_load_volume.argtypes = [C.c_char_p, _u32p, _f32p, _f32p, C.POINTER(u32), _u32p, u32, _f32p, C.c_uint64]
    if hasattr(R, "ref_nelder_mead_step"):
        R.ref_nelder_mead_step.argtypes = [_f32p, REF_OBJECTIVE, _vp, C.POINTER(u32), C.POINTER(f32)]
    _REFGEOM = R
    return R


def set_threads(n=0):
    """Sets the OpenMP team size of the oracle's loops (0: keep) and returns the team size a parallel region gets."""
    return int(lib().orc_set_threads(int(n)))


def ref():
    """The reference's own acc code (None when oracle/_ref was not built)."""
    global _REF
    if _REF is not None:
        return _REF
    build()
    path = os.path.join(_HERE, "_ref", "libref_acc.so")
    if not os.path.exists(path):
        return None
    R = C.CDLL(path)
    u32, u64, i32 = C.c_uint32, C.c_uint64, C.c_int
    R.ref_bvh_build.restype = _vp
    R.ref_bvh_build.argtypes = [_f32p, u32, _u32p, u32, i32]
    R.ref_bvh_num_nodes.restype = u32
    R.ref_bvh_num_nodes.argtypes = [_vp]
    R.ref_bvh_num_tris.restype = u32
    R.ref_bvh_num_tris.argtypes = [_vp]
    R.ref_bvh_export.argtypes = [_vp, _u32p, _f32p, _f32p, _u32p]
    R.ref_bvh_intersect.argtypes = [_vp, _f32p, u64, _opt(_u8p), _opt(_u32p), _opt(_f32p)]
    R.ref_bvh_closest_point.argtypes = [_vp, _f32p, u64, C.c_float, _u8p, _u32p, _f32p]
    R.ref_bvh_destroy.argtypes = [_vp]
    R.ref_kd_build.restype = _vp
    R.ref_kd_build.argtypes = [_f32p, u32, i32]
    R.ref_kd_num_nodes.restype = u32
    R.ref_kd_num_nodes.argtypes = [_vp]
    R.ref_kd_export.argtypes = [_vp, _u32p, _f32p]
    R.ref_kd_find_nn.argtypes = [_vp, _f32p, u64, _u32p, _opt(_f32p)]
    R.ref_kd_destroy.argtypes = [_vp]
    _REF = R
    return R


_REFCU = {}


def refcuda(kind="exact"):
    """The reference's own CUDA kernels compiled for sm_100a (None when oracle/_ref holds no such build or
    there is no GPU).  kind: "exact" (IEEE, no FMA contraction) or "fast" (the reference's nvcc flags)."""
    if kind in _REFCU:
        return _REFCU[kind]
    build()
    path = os.path.join(_HERE, "_ref", "libref_cuda_%s.so" % kind)
    if not os.path.exists(path):
        return None
    R = C.CDLL(path)
    if R.refcu_device_count() <= 0:
        return None
    u32, i32, f32 = C.c_uint32, C.c_int, C.c_float
    for name in ("refcu_bvh_create", "refcu_kd_create", "refcu_cloud_create", "refcu_obs_create"):
        getattr(R, name).restype = _vp
    R.refcu_bvh_create.argtypes = [_f32p, u32, _u32p, u32, i32]
    R.refcu_kd_create.argtypes = [_f32p, u32]
    R.refcu_cloud_create.argtypes = [u32, _f32p, _f32p, _opt(_f32p), _opt(_f32p)]
    R.refcu_cloud_download.argtypes = [_vp, _opt(_f32p), _opt(_f32p)]
    R.refcu_obs_create.argtypes = [u32, u32]
    R.refcu_obs_clear.argtypes = [_vp]
    R.refcu_obs_download.argtypes = [_vp, _opt(_u32p), _opt(_f32p)]
    for name in ("refcu_bvh_destroy", "refcu_kd_destroy", "refcu_cloud_destroy", "refcu_obs_destroy",
                 "refcu_process_observation_rays"):
        getattr(R, name).argtypes = [_vp]
    R.refcu_configure_heuristic.argtypes = [f32, f32, f32, f32]
    R.refcu_update_observation_rays.argtypes = [i32, _f32p, u32, f32, i32, i32, _vp, _vp, _vp]
    R.refcu_evaluate_observation_rays.argtypes = [_vp, _f32p]
    R.refcu_calculate_func_recons.argtypes = [_f32p, u32, f32, _opt(_f32p), _opt(_f32p)]
    R.refcu_populate_spherical_histogram.argtypes = [_f32p, f32, f32, _vp, _vp, _vp, _vp, _opt(_f32p), u32, _f32p]
    R.refcu_evaluate_spherical_histogram.argtypes = [_f32p, i32, i32, _vp, _f32p, u32, _f32p]
    R.refcu_estimate_capture_difficulty.argtypes = [f32, _vp, u32, _vp, _vp]
    R.refcu_time_evaluate.restype = f32
    R.refcu_time_evaluate.argtypes = [_f32p, u32, f32, i32, i32, f32, _vp, _vp, _vp]
    _REFCU[kind] = R
    return R


class RefCuda:
    """One scene held by the reference's own containers on the GPU (cacc::BVHTree / PointCloud / VectorArray /
    KDTree), driven through the reference's kernels the way its apps drive them."""

    def __init__(self, verts, faces, points, normals, values=None, qualities=None, sphere_verts=None,
                 max_rows=32, kind="exact", threads=0):
        self.R = refcuda(kind)
        if self.R is None:
            raise RuntimeError("oracle/_ref/libref_cuda_%s.so not available (no build, or no GPU)" % kind)
        verts = _c(verts, np.float32).reshape(-1, 3)
        faces = _c(faces, np.uint32).reshape(-1, 3)
        self.n = len(points)
        self.max_rows = max_rows
        self.bvh = self.R.refcu_bvh_create(verts, verts.shape[0], faces, faces.shape[0], threads)
        self.cloud = self.R.refcu_cloud_create(self.n, _c(points, np.float32), _c(normals, np.float32),
                                               None if values is None else _c(values, np.float32),
                                               None if qualities is None else _c(qualities, np.float32))
        self.obs = self.R.refcu_obs_create(self.n, max_rows)
        self.kd, self.S = None, 0
        if sphere_verts is not None:
            sv = _c(sphere_verts, np.float32).reshape(-1, 3)
            self.kd, self.S = self.R.refcu_kd_create(sv, sv.shape[0]), sv.shape[0]

    def configure_heuristic(self, m_k, m_x0, t_k, t_x0):
        self.R.refcu_configure_heuristic(m_k, m_x0, t_k, t_x0)

    def clear(self):
        self.R.refcu_obs_clear(self.obs)

    def update_observation_rays(self, populate, views, max_distance, width, height):
        views = _c(views, np.float32).reshape(-1, VIEW_FLOATS)
        self.R.refcu_update_observation_rays(1 if populate else 0, views, views.shape[0], max_distance, width, height,
                                             self.bvh, self.cloud, self.obs)

    def process_observation_rays(self):
        self.R.refcu_process_observation_rays(self.obs)

    def evaluate_observation_rays(self):
        out = np.empty(self.n, np.float32)
        self.R.refcu_evaluate_observation_rays(self.obs, out)
        return out

    def download(self):
        """(num_rows[N] uint32, rows[max_rows, N, 4] float32)"""
        num_rows = np.empty(self.n, np.uint32)
        rows = np.empty((self.max_rows, self.n, 4), np.float32)
        self.R.refcu_obs_download(self.obs, num_rows, rows)
        return num_rows, rows

    def calculate_func_recons(self, recons, target):
        recons = _c(recons, np.float32)
        w = np.empty_like(recons)
        out3 = np.zeros(3, np.float32)
        self.R.refcu_calculate_func_recons(recons, recons.shape[0], target, w, out3)
        return w, out3

    def populate_spherical_histogram(self, view_pos, max_distance, target_recon=0.0, recons=None):
        hist = np.empty(self.S, np.float32)
        self.R.refcu_populate_spherical_histogram(_c(view_pos, np.float32), max_distance, target_recon, self.bvh,
                                                  self.cloud, self.kd, self.obs,
                                                  None if recons is None else _c(recons, np.float32), self.S, hist)
        return hist

    def evaluate_spherical_histogram(self, calib, width, height, hist):
        img = np.empty((45, 128), np.float32)
        self.R.refcu_evaluate_spherical_histogram(_c(calib, np.float32).reshape(9), width, height, self.kd,
                                                  _c(hist, np.float32), self.S, img)
        return img

    def estimate_capture_difficulty(self, max_distance, mesh_size):
        self.R.refcu_estimate_capture_difficulty(max_distance, self.bvh, mesh_size, self.kd, self.cloud)
        values, qualities = np.empty(self.n, np.float32), np.empty(self.n, np.float32)
        self.R.refcu_cloud_download(self.cloud, values, qualities)
        return values, qualities

    def time_evaluate(self, views, max_distance, width, height, target_recon):
        views = _c(views, np.float32).reshape(-1, VIEW_FLOATS)
        return float(self.R.refcu_time_evaluate(views, views.shape[0], max_distance, width, height, target_recon,
                                                self.bvh, self.cloud, self.obs))

    def __del__(self):
        try:
            self.R.refcu_obs_destroy(self.obs)
            self.R.refcu_cloud_destroy(self.cloud)
            self.R.refcu_bvh_destroy(self.bvh)
            if self.kd:
                self.R.refcu_kd_destroy(self.kd)
        except Exception:
            pass


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


class TreeArrays:
    """Flat export of a binary BVH in the reference's layout
    (libs/cacc/bvh_tree.h:31-42, :86-93): nodes[n,4] = first,last,left,right."""

    def __init__(self, nodes, aabbs, tris, indices):
        self.nodes, self.aabbs, self.tris, self.indices = nodes, aabbs, tris, indices

    @property
    def num_nodes(self):
        return self.nodes.shape[0]

    @property
    def num_tris(self):
        return self.indices.shape[0]


class BVH:
    """Oracle BVH (our restatement of acc::BVHTree)."""

    def __init__(self, handle):
        self.h = handle

    @classmethod
    def build(cls, verts, faces, threads=0):
        verts = _c(verts, np.float32).reshape(-1, 3)
        faces = _c(faces, np.uint32).reshape(-1, 3)
        threads = threads or os.cpu_count() or 1
        return cls(lib().orc_bvh_build(verts, verts.shape[0], faces, faces.shape[0], threads))

    @classmethod
    def from_arrays(cls, t):
        return cls(lib().orc_bvh_from_arrays(_c(t.nodes, np.uint32), _c(t.aabbs, np.float32),
                                             _c(t.tris, np.float32), _c(t.indices, np.uint32),
                                             t.num_nodes, t.num_tris))

    def export(self):
        L = lib()
        n, t = L.orc_bvh_num_nodes(self.h), L.orc_bvh_num_tris(self.h)
        nodes = np.empty((n, 4), np.uint32)
        aabbs = np.empty((n, 6), np.float32)
        tris = np.empty((t, 9), np.float32)
        idx = np.empty(t, np.uint32)
        L.orc_bvh_export(self.h, nodes, aabbs, tris, idx)
        return TreeArrays(nodes, aabbs, tris, idx)

    def closest_point(self, points, max_dist=float("inf")):
        """acc::BVHTree::closest_point (libs/acc/bvh_tree.h:492-548): found, leaf-order triangle index, point."""
        pts = _c(points, np.float32).reshape(-1, 3)
        n = pts.shape[0]
        found, idx, cp = np.empty(n, np.uint8), np.empty(n, np.uint32), np.empty((n, 3), np.float32)
        lib().orc_bvh_closest_point(self.h, pts, n, max_dist, found, idx, cp)
        return found, idx, cp

    def intersect(self, rays, any_hit=True, counters=False):
        rays = _c(rays, np.float32).reshape(-1, 8)
        n = rays.shape[0]
        hit = np.empty(n, np.uint8)
        face = np.empty(n, np.uint32)
        t = np.empty(n, np.float32)
        cnt = np.zeros(5, np.uint64) if counters else None
        lib().orc_bvh_intersect(self.h, rays, n, 1 if any_hit else 0, hit, face, t, cnt)
        return (hit, face, t, cnt) if counters else (hit, face, t)

    def __del__(self):
        if getattr(self, "h", None) and _LIB is not None:
            _LIB.orc_bvh_destroy(self.h)
            self.h = None


class RefBVH:
    """The reference's own acc::BVHTree (oracle/_ref)."""

    def __init__(self, verts, faces, threads=0):
        R = ref()
        if R is None:
            raise RuntimeError("oracle/_ref/libref_acc.so not built")
        verts = _c(verts, np.float32).reshape(-1, 3)
        faces = _c(faces, np.uint32).reshape(-1, 3)
        self.h = R.ref_bvh_build(verts, verts.shape[0], faces, faces.shape[0], threads or os.cpu_count() or 1)

    def export(self):
        R = ref()
        n, t = R.ref_bvh_num_nodes(self.h), R.ref_bvh_num_tris(self.h)
        nodes = np.empty((n, 4), np.uint32)
        aabbs = np.empty((n, 6), np.float32)
        tris = np.empty((t, 9), np.float32)
        idx = np.empty(t, np.uint32)
        R.ref_bvh_export(self.h, nodes, aabbs, tris, idx)
        return TreeArrays(nodes, aabbs, tris, idx)

    def closest_point(self, points, max_dist=float("inf")):
        pts = _c(points, np.float32).reshape(-1, 3)
        n = pts.shape[0]
        found, idx, cp = np.empty(n, np.uint8), np.empty(n, np.uint32), np.empty((n, 3), np.float32)
        ref().ref_bvh_closest_point(self.h, pts, n, max_dist, found, idx, cp)
        return found, idx, cp

    def intersect(self, rays):
        rays = _c(rays, np.float32).reshape(-1, 8)
        n = rays.shape[0]
        hit = np.empty(n, np.uint8)
        face = np.empty(n, np.uint32)
        t = np.empty(n, np.float32)
        ref().ref_bvh_intersect(self.h, rays, n, hit, face, t)
        return hit, face, t

    def __del__(self):
        if getattr(self, "h", None) and _REF is not None:
            _REF.ref_bvh_destroy(self.h)
            self.h = None


class KD:
    def __init__(self, handle):
        self.h = handle

    @classmethod
    def build(cls, verts):
        verts = _c(verts, np.float32).reshape(-1, 3)
        return cls(lib().orc_kd_build(verts, verts.shape[0]))

    @classmethod
    def from_arrays(cls, nodes, verts):
        nodes = _c(nodes, np.uint32).reshape(-1, 4)
        verts = _c(verts, np.float32).reshape(-1, 3)
        return cls(lib().orc_kd_from_arrays(nodes, verts, nodes.shape[0], verts.shape[0]))

    def export(self):
        L = lib()
        n = L.orc_kd_num_nodes(self.h)
        nodes = np.empty((n, 4), np.uint32)
        verts = np.empty((n, 3), np.float32)
        L.orc_kd_export(self.h, nodes, verts)
        return nodes, verts

    def find_nn(self, q):
        q = _c(q, np.float32).reshape(-1, 3)
        idx = np.empty(q.shape[0], np.uint32)
        dist = np.empty(q.shape[0], np.float32)
        lib().orc_kd_find_nn(self.h, q, q.shape[0], idx, dist)
        return idx, dist

    def __del__(self):
        if getattr(self, "h", None) and _LIB is not None:
            _LIB.orc_kd_destroy(self.h)
            self.h = None


def sphere_vertices(radius=1.0, subdivisions=3):
    n = lib().orc_sphere_vertices(radius, subdivisions, None, 0)
    out = np.empty((n, 3), np.float32)
    lib().orc_sphere_vertices(radius, subdivisions, out, n)
    return out


def pack_views(view_pos, w2c, calib):
    """[V,3], [V,4,4], [V,3,3] -> [V,28] float32."""
    V = len(view_pos)
    out = np.empty((V, VIEW_FLOATS), np.float32)
    out[:, 0:3] = np.asarray(view_pos, np.float32).reshape(V, 3)
    out[:, 3:19] = np.asarray(w2c, np.float32).reshape(V, 16)
    out[:, 19:28] = np.asarray(calib, np.float32).reshape(V, 9)
    return out


class ObsRays:
    """Host mirror of cacc::VectorArray<Vec3f> (libs/cacc/vector_array.h:23-50):
    data[max_rows, stride, 4] float32, num_rows[N] uint32."""

    def __init__(self, n, max_rows=32, stride=None):
        self.n, self.max_rows = n, max_rows
        self.stride = stride or n
        self.data = np.zeros((max_rows, self.stride, 4), np.float32)
        self.num_rows = np.zeros(n, np.uint32)


def update_observation_rays(bvh, populate, views, max_distance, width, height, verts, normals, obs,
                            want_state=False, want_counters=False):
    views = _c(views, np.float32).reshape(-1, VIEW_FLOATS)
    verts = _c(verts, np.float32).reshape(-1, 3)
    normals = _c(normals, np.float32).reshape(-1, 3)
    n = verts.shape[0]
    state = np.zeros((views.shape[0], n), np.uint8) if want_state else None
    cnt = np.zeros(6, np.uint64) if want_counters else None
    lib().orc_update_observation_rays(bvh.h, 1 if populate else 0, views, views.shape[0], max_distance,
                                      width, height, verts, normals, n, obs.data, obs.num_rows,
                                      obs.max_rows, obs.stride, state, cnt)
    return state, cnt


def process_observation_rays(obs):
    lib().orc_process_observation_rays(obs.data, obs.num_rows, obs.max_rows, obs.stride, obs.n)


def evaluate_observation_rays(obs, heur=(8.0, 4.0, 32.0, 16.0)):
    recons = np.empty(obs.n, np.float32)
    lib().orc_evaluate_observation_rays(obs.data, obs.num_rows, obs.max_rows, obs.stride, obs.n,
                                        np.asarray(heur, np.float32), recons)
    return recons


def calculate_func_recons(recons, target):
    out = np.empty_like(recons)
    lib().orc_calculate_func_recons(_c(recons, np.float32), target, out, recons.shape[0])
    return out


def reduce(x):
    out = np.zeros(3, np.float64)
    lib().orc_reduce(_c(x, np.float32), x.shape[0], out)
    return out


def populate_spherical_histogram(bvh, kd, view_pos, max_distance, verts, normals, values, qualities, S):
    n = len(verts)
    hist = np.zeros(S, np.float64)
    bins = np.empty(n, np.uint32)
    scores = np.empty(n, np.float32)
    lib().orc_populate_spherical_histogram(bvh.h, kd.h, _c(view_pos, np.float32), max_distance,
                                           _c(verts, np.float32), _c(normals, np.float32),
                                           _c(values, np.float32), _c(qualities, np.float32), n,
                                           hist, bins, scores)
    return hist, bins, scores


def populate_spherical_histogram_recon(bvh, kd, view_pos, max_distance, target_recon, verts, normals,
                                       obs, recons, S, heur=(8.0, 4.0, 32.0, 16.0)):
    n = len(verts)
    hist = np.zeros(S, np.float64)
    bins = np.empty(n, np.uint32)
    deltas = np.empty(n, np.float32)
    lib().orc_populate_spherical_histogram_recon(
        bvh.h, kd.h, _c(view_pos, np.float32), max_distance, target_recon,
        _c(verts, np.float32), _c(normals, np.float32), n, obs.data, obs.num_rows, obs.max_rows,
        obs.stride, _c(recons, np.float32), np.asarray(heur, np.float32), hist, bins, deltas)
    return hist, bins, deltas


def evaluate_spherical_histogram(calib, width, height, sphere_verts, sphere_hist, hist_w=128, hist_h=45):
    sv = _c(sphere_verts, np.float32).reshape(-1, 3)
    out = np.empty((hist_h, hist_w), np.float32)
    lib().orc_evaluate_spherical_histogram(_c(calib, np.float32).reshape(9), width, height, sv, sv.shape[0],
                                           _c(sphere_hist, np.float32), out, hist_w, hist_h)
    return out


def estimate_capture_difficulty(bvh, max_distance, mesh_size, sphere_verts, verts, normals):
    sv = _c(sphere_verts, np.float32).reshape(-1, 3)
    n = len(verts)
    values = np.empty(n, np.float32)
    qualities = np.empty(n, np.float32)
    lib().orc_estimate_capture_difficulty(bvh.h, max_distance, mesh_size, sv, sv.shape[0],
                                          _c(verts, np.float32), _c(normals, np.float32), n, values, qualities)
    return values, qualities


def sample_proxy_cloud(verts, faces, samples):
    """apps/generate_proxy_cloud/generate_proxy_cloud.cpp:84-184 -> (points [n,3], normals [n,3]), Z-order sorted."""
    verts = _c(verts, np.float32).reshape(-1, 3)
    faces = _c(faces, np.uint32).reshape(-1, 3)
    n = lib().orc_sample_proxy_cloud(verts, faces, faces.shape[0], samples, None, None, 0)
    pts, nrm = np.empty((n, 3), np.float32), np.empty((n, 3), np.float32)
    if n:
        lib().orc_sample_proxy_cloud(verts, faces, faces.shape[0], samples, pts, nrm, n)
    return pts, nrm
