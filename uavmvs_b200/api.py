"""Host-side mirror of the reference's cacc:: / libs/eval interface over the C ABI.

Class and function names follow the reference so that callers (and the parity
tests) read like its apps:

    cacc::BVHTree<DEVICE>::create(acc_tree)        -> BVHTree.create(nodes, aabbs, tris, indices)
    cacc::KDTree<3,DEVICE>::create(acc_kd)         -> KDTree.create(nodes, verts)
    cacc::PointCloud<DEVICE>::create<HOST>(cloud)  -> PointCloud.create(vertices, normals, ...)
    cacc::VectorArray<Vec3f,DEVICE>::create(n, 32) -> VectorArray.create(n, 32)
    cacc::Array<float,DEVICE>::create(n)           -> Array.create(n)
    cacc::Image<float,DEVICE>::create(w, h)        -> Image.create(w, h)
    update_observation_rays<<<...>>>(...)          -> update_observation_rays(...)   (etc.)

There is no CPU fallback: loading this module without the compiled CUDA library,
or calling anything without a CUDA device, raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.environ.get("UAVMVS_B200_LIB") or os.path.join(_HERE, "libuavmvs_b200.so")   # override: dev variants
VIEW_FLOATS = 28
NAI = 0xFFFFFFFF


class UavmvsError(RuntimeError):
    pass


def _load():
    if not os.path.exists(_SO):
        raise UavmvsError(
            f"{_SO} is missing: build it with `python -m uavmvs_b200.build` "
            "(there is no CPU fallback for the evaluation path)")
    return C.CDLL(_SO)


_lib = _load()
_lib.uavmvs_last_error.restype = C.c_char_p

_vp, _u32, _u64, _i32, _f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_float


def _check(rc):
    if rc != 0:
        raise UavmvsError(_lib.uavmvs_last_error().decode())


def _f(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a if shape is None else a.reshape(shape)


def _u(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def _p(a):
    return None if a is None else a.ctypes.data_as(_vp)


def _stream_ptr(stream):
    """Accepts None, an int (cudaStream_t) or a torch.cuda.Stream."""
    if stream is None:
        return _vp(0)
    if hasattr(stream, "cuda_stream"):
        return _vp(stream.cuda_stream)
    return _vp(int(stream))


def set_device(device):
    _check(_lib.uavmvs_set_device(_i32(device)))


def device_count():
    n = _i32(0)
    _check(_lib.uavmvs_device_count(C.byref(n)))
    return n.value


def synchronize(stream=None):
    _check(_lib.uavmvs_synchronize(_stream_ptr(stream)))


def configure_heuristic(m_k=8.0, m_x0=4.0, t_k=32.0, t_x0=16.0):
    """libs/eval/kernels.h:93"""
    _check(_lib.uavmvs_configure_heuristic(_f32(m_k), _f32(m_x0), _f32(t_k), _f32(t_x0)))


class _Handle:
    _destroy = None

    def __init__(self, ptr):
        self.ptr = ptr

    def free(self):
        if getattr(self, "ptr", None):
            getattr(_lib, self._destroy)(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class BVHTree(_Handle):
    """cacc::BVHTree<DEVICE> (libs/cacc/bvh_tree.h)."""
    _destroy = "uavmvs_bvh_destroy"

    @classmethod
    def create(cls, nodes, aabbs, tris, indices):
        """From an acc-built tree in the layout convert<HOST> produces
        (libs/cacc/bvh_tree.cu:17-48)."""
        nodes, aabbs, tris, indices = _u(nodes).reshape(-1, 4), _f(aabbs, (-1, 6)), _f(tris, (-1, 9)), _u(indices)
        h = _vp()
        _check(_lib.uavmvs_bvh_create_from_arrays(_p(nodes), _p(aabbs), _p(tris), _p(indices),
                                                  _u32(nodes.shape[0]), _u32(indices.shape[0]), C.byref(h)))
        return cls(h)

    @classmethod
    def build_lbvh(cls, verts, faces):
        """GPU LBVH build from an indexed mesh."""
        verts, faces = _f(verts, (-1, 3)), _u(faces).reshape(-1, 3)
        h = _vp()
        _check(_lib.uavmvs_bvh_build_lbvh(_p(verts), _u32(verts.shape[0]), _p(faces), _u32(faces.shape[0]),
                                          C.byref(h)))
        return cls(h)

    @classmethod
    def build_ploc(cls, verts, faces):
        """GPU build by parallel locally-ordered clustering (SAH-like quality)."""
        verts, faces = _f(verts, (-1, 3)), _u(faces).reshape(-1, 3)
        h = _vp()
        _check(_lib.uavmvs_bvh_build_ploc(_p(verts), _u32(verts.shape[0]), _p(faces), _u32(faces.shape[0]),
                                          C.byref(h)))
        return cls(h)

    @classmethod
    def build_sah(cls, verts, faces):
        """GPU full-sweep SAH build (the reference builder's cost model, libs/acc/bvh_tree.h:277-333)."""
        verts, faces = _f(verts, (-1, 3)), _u(faces).reshape(-1, 3)
        h = _vp()
        _check(_lib.uavmvs_bvh_build_sah(_p(verts), _u32(verts.shape[0]), _p(faces), _u32(faces.shape[0]),
                                         C.byref(h)))
        return cls(h)

    def sizes(self):
        n, t = _u32(0), _u32(0)
        _check(_lib.uavmvs_bvh_sizes(self.ptr, C.byref(n), C.byref(t)))
        return n.value, t.value

    def export(self):
        n, t = self.sizes()
        nodes, aabbs = np.empty((n, 4), np.uint32), np.empty((n, 6), np.float32)
        tris, indices = np.empty((t, 9), np.float32), np.empty(t, np.uint32)
        _check(_lib.uavmvs_bvh_export(self.ptr, _p(nodes), _p(aabbs), _p(tris), _p(indices)))
        return nodes, aabbs, tris, indices

    def closest_point(self, points, max_dist=float("inf"), stream=None):
        """acc::BVHTree::closest_point for a batch: (found u8[n], leaf-order triangle index u32[n], point f32[n,3])."""
        pts = _f(points).reshape(-1, 3)
        n = pts.shape[0]
        found, idx, cp = np.empty(n, np.uint8), np.empty(n, np.uint32), np.empty((n, 3), np.float32)
        _check(_lib.uavmvs_bvh_closest_point(_stream_ptr(stream), self.ptr, _p(pts), C.c_uint64(n), _f32(max_dist),
                                             _p(found), _p(idx), _p(cp)))
        return found, idx, cp

    def trace(self, rays, any_hit=True, stream=None):
        """cacc::tracing::trace over a batch of host rays [n,8]."""
        rays = _f(rays, (-1, 8))
        n = rays.shape[0]
        hit, face, t = np.zeros(n, np.uint8), np.zeros(n, np.uint32), np.zeros(n, np.float32)
        _check(_lib.uavmvs_bvh_trace(_stream_ptr(stream), self.ptr, _p(rays), _u64(n), _i32(1 if any_hit else 0),
                                     _p(hit), _p(face), _p(t)))
        return hit, face, t


class KDTree(_Handle):
    """cacc::KDTree<3, DEVICE> (libs/cacc/kd_tree.h)."""
    _destroy = "uavmvs_kdtree_destroy"

    @classmethod
    def create(cls, nodes, verts):
        nodes, verts = _u(nodes).reshape(-1, 4), _f(verts, (-1, 3))
        h = _vp()
        _check(_lib.uavmvs_kdtree_create(_p(nodes), _p(verts), _u32(nodes.shape[0]), C.byref(h)))
        obj = cls(h)
        obj.num_verts = nodes.shape[0]
        return obj


    @classmethod
    def build(cls, verts):
        """Median-split build on the host (acc::KDTree<3>'s rule) + upload, for callers without libs/acc."""
        verts = _f(verts, (-1, 3))
        h = _vp()
        _check(_lib.uavmvs_kdtree_build(_p(verts), _u32(verts.shape[0]), C.byref(h)))
        obj = cls(h)
        obj.num_verts = verts.shape[0]
        return obj

    def find_nn(self, queries, certified_lookup=False, stream=None):
        """cacc::nnsearch::find_nn<3> for a batch of queries [n,3] -> vertex indices."""
        q = _f(queries).reshape(-1, 3)
        out = np.empty(q.shape[0], np.uint32)
        _check(_lib.uavmvs_kdtree_find_nn(_stream_ptr(stream), self.ptr, _p(q), _u64(q.shape[0]),
                                          _i32(1 if certified_lookup else 0), _p(out)))
        return out


class PointCloud(_Handle):
    """cacc::PointCloud<DEVICE> (libs/cacc/point_cloud.h)."""
    _destroy = "uavmvs_cloud_destroy"

    @classmethod
    def create(cls, vertices, normals, values=None, qualities=None):
        vertices, normals = _f(vertices, (-1, 3)), _f(normals, (-1, 3))
        n = vertices.shape[0]
        values = None if values is None else _f(values)
        qualities = None if qualities is None else _f(qualities)
        h = _vp()
        _check(_lib.uavmvs_cloud_create(_p(vertices), _p(normals), _p(values), _p(qualities), _u32(n), C.byref(h)))
        obj = cls(h)
        obj.num_vertices = n
        return obj

    def download_attributes(self):
        values, qualities = np.empty(self.num_vertices, np.float32), np.empty(self.num_vertices, np.float32)
        _check(_lib.uavmvs_cloud_download_attributes(self.ptr, _p(values), _p(qualities)))
        return values, qualities

    def upload_attributes(self, values=None, qualities=None):
        values = None if values is None else _f(values)
        qualities = None if qualities is None else _f(qualities)
        _check(_lib.uavmvs_cloud_upload_attributes(self.ptr, _p(values), _p(qualities)))


class VectorArray(_Handle):
    """cacc::VectorArray<cacc::Vec3f, DEVICE> (libs/cacc/vector_array.h)."""
    _destroy = "uavmvs_obs_destroy"

    @classmethod
    def create(cls, num_cols, max_rows=32):
        h = _vp()
        _check(_lib.uavmvs_obs_create(_u32(num_cols), _u32(max_rows), C.byref(h)))
        obj = cls(h)
        obj.num_cols, obj.max_rows = num_cols, max_rows
        return obj

    def configure_heuristic(self, m_k=8.0, m_x0=4.0, t_k=32.0, t_x0=16.0):
        """Heuristic constants of this observation array only (the process-wide ones stay as they are)."""
        _check(_lib.uavmvs_obs_configure_heuristic(self.ptr, _f32(m_k), _f32(m_x0), _f32(t_k), _f32(t_x0)))

    def null(self, stream=None):
        _check(_lib.uavmvs_obs_clear(_stream_ptr(stream), self.ptr))

    def mark_dirty(self, stream=None):
        _check(_lib.uavmvs_obs_mark_dirty(_stream_ptr(stream), self.ptr))

    def download(self, data=True, rows_out=None):
        d = np.empty((self.max_rows, self.num_cols, 4), np.float32) if data else None
        r = np.empty(self.num_cols, np.uint32) if rows_out is None else rows_out
        assert r.dtype == np.uint32 and r.size == self.num_cols
        _check(_lib.uavmvs_obs_download(self.ptr, _p(d), _p(r)))
        return d, r

    def upload(self, data=None, num_rows=None):
        d = None if data is None else _f(data, (self.max_rows, self.num_cols, 4))
        r = None if num_rows is None else _u(num_rows)
        _check(_lib.uavmvs_obs_upload(self.ptr, _p(d), _p(r)))


class Array(_Handle):
    """cacc::Array<float, DEVICE> (libs/cacc/array.h)."""
    _destroy = "uavmvs_array_destroy"

    @classmethod
    def create(cls, n):
        h = _vp()
        _check(_lib.uavmvs_array_create(_u32(n), C.byref(h)))
        obj = cls(h)
        obj.n = n
        return obj

    def null(self, stream=None):
        _check(_lib.uavmvs_array_null(_stream_ptr(stream), self.ptr))

    def download(self, out=None):
        """`out`: optional host float32 array of n elements (e.g. a view of pinned memory)."""
        if out is None:
            out = np.empty(self.n, np.float32)
        assert out.dtype == np.float32 and out.size == self.n and out.flags["C_CONTIGUOUS"]
        _check(_lib.uavmvs_array_download(self.ptr, _p(out)))
        return out

    def upload(self, host):
        host = _f(host)
        assert host.shape[0] == self.n
        _check(_lib.uavmvs_array_upload(self.ptr, _p(host)))


class Image(_Handle):
    """cacc::Image<float, DEVICE> (libs/cacc/image.h)."""
    _destroy = "uavmvs_image_destroy"

    @classmethod
    def create(cls, width, height):
        h = _vp()
        _check(_lib.uavmvs_image_create(_u32(width), _u32(height), C.byref(h)))
        obj = cls(h)
        obj.width, obj.height = width, height
        return obj

    def download(self):
        out = np.empty((self.height, self.width), np.float32)
        _check(_lib.uavmvs_image_download(self.ptr, _p(out)))
        return out


# ------------------------------------------------------------------ kernels.h
def update_observation_rays(populate, view_pos, max_distance, w2c, calib, width, height,
                            bvh_tree, cloud, obs_rays, stream=None):
    """libs/eval/kernels.h:26-35 (one view)."""
    vp, W, K = _f(view_pos, 3), _f(w2c, 16), _f(calib, 9)
    _check(_lib.uavmvs_update_observation_rays(_stream_ptr(stream), _i32(1 if populate else 0), _p(vp),
                                               _f32(max_distance), _p(W), _p(K), _i32(width), _i32(height),
                                               bvh_tree.ptr, cloud.ptr, obs_rays.ptr))


def update_observation_rays_batch(populate, views, max_distance, width, height, bvh_tree, cloud, obs_rays,
                                  stream=None):
    """The launch loop of apps/evaluate_trajectory/evaluate_trajectory.cu:154-164 as one call;
    views [V,28] = pos, w2c, calib (host)."""
    if isinstance(views, DeviceViews):
        _check(_lib.uavmvs_update_observation_rays_batch_device(_stream_ptr(stream), _i32(1 if populate else 0), views.ptr,
                                                                _u32(views.num_views), _f32(max_distance), _i32(width),
                                                                _i32(height), bvh_tree.ptr, cloud.ptr, obs_rays.ptr))
        return
    views = _f(views, (-1, VIEW_FLOATS))
    _check(_lib.uavmvs_update_observation_rays_batch(_stream_ptr(stream), _i32(1 if populate else 0), _p(views),
                                                     _u32(views.shape[0]), _f32(max_distance), _i32(width),
                                                     _i32(height), bvh_tree.ptr, cloud.ptr, obs_rays.ptr))


def process_observation_rays(obs_rays, stream=None):
    """libs/eval/kernels.h:37-39"""
    _check(_lib.uavmvs_process_observation_rays(_stream_ptr(stream), obs_rays.ptr))


def evaluate_observation_rays(obs_rays, recons, stream=None):
    """libs/eval/kernels.h:41-44"""
    _check(_lib.uavmvs_evaluate_observation_rays(_stream_ptr(stream), obs_rays.ptr, recons.ptr))


def reevaluate_observation_rays(obs_rays, recons, target_recon=0.0, wrecons=None, stream=None):
    """K2 + K3 (+ K8) on the 32-column blocks update_observation_rays touched since the last call
    (the re-evaluation of apps/optimize_trajectory/optimize_trajectory.cu:313-345, :499-529)."""
    _check(_lib.uavmvs_reevaluate_observation_rays(_stream_ptr(stream), obs_rays.ptr, recons.ptr, _f32(target_recon),
                                                   None if wrecons is None else wrecons.ptr))


def calculate_func_recons(recons, target_recon, wrecons, stream=None):
    """libs/eval/kernels.h:86-91"""
    _check(_lib.uavmvs_calculate_func_recons(_stream_ptr(stream), recons.ptr, _f32(target_recon), wrecons.ptr))


def reduce(array, stream=None):
    """cacc::reduction::{sum,min,max} (libs/cacc/reduction.h:132-162) -> (sum, min, max)."""
    out = (C.c_float * 3)()
    _check(_lib.uavmvs_reduce(_stream_ptr(stream), array.ptr, out))
    return float(out[0]), float(out[1]), float(out[2])


REDUCE_NOALLOC_MAX_N = 1 << 24      # UAVMVS_REDUCE_NOALLOC_MAX_N


def reduce_noalloc(array, out3, stream=None):
    """uavmvs_reduce_noalloc: K9 into `out3` (an Array of >= 3 floats created beforehand) without any allocation or
    synchronisation; no counterpart in the reference.  Read the result with out3.download()[:3] after the stream is done.
    Calls on one device share a static scratch array: do not overlap them."""
    _check(_lib.uavmvs_reduce_noalloc(_stream_ptr(stream), array.ptr, out3.ptr))


class DeviceViews:
    """A trajectory resident on the device ([V,28] floats) for repeated evaluation (uavmvs_views_upload)."""

    def __init__(self, views, stream=None):
        v = _f(views, (-1, VIEW_FLOATS))
        self.num_views = int(v.shape[0])
        d = C.POINTER(C.c_float)()
        _check(_lib.uavmvs_views_upload(_stream_ptr(stream), _p(v), _u32(self.num_views), C.byref(d)))
        self._d = d

    @property
    def ptr(self):
        return C.cast(self._d, _vp)

    def update(self, views, stream=None):
        """Overwrites the views (same count) -- a stream-ordered H2D copy, no allocation."""
        v = _f(views, (-1, VIEW_FLOATS))
        assert v.shape[0] == self.num_views
        self._host = v                                      # keep the source alive until the copy ran
        _check(_lib.uavmvs_views_update(_stream_ptr(stream), self._d, _p(v), _u32(self.num_views)))

    def download(self, stream=None):
        out = np.empty((self.num_views, VIEW_FLOATS), np.float32)
        _check(_lib.uavmvs_views_download(_stream_ptr(stream), self._d, _u32(self.num_views), _p(out)))
        return out

    def free(self):
        if getattr(self, "_d", None):
            _lib.uavmvs_views_free(self._d)
            self._d = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Comm(_Handle):
    """One rank of the point-sharded evaluation: an ncclComm_t behind the C ABI (include/uavmvs_b200.h)."""
    _destroy = "uavmvs_comm_destroy"
    ID_BYTES = 128

    @staticmethod
    def unique_id():
        """ncclGetUniqueId: call on one rank, ship the 128 bytes to the others."""
        buf = (C.c_uint8 * Comm.ID_BYTES)()
        _check(_lib.uavmvs_comm_unique_id(buf))
        return bytes(buf)

    @classmethod
    def create(cls, unique_id, world, rank):
        """ncclCommInitRank on the current device (collective)."""
        buf = (C.c_uint8 * cls.ID_BYTES).from_buffer_copy(unique_id)
        h = _vp()
        _check(_lib.uavmvs_comm_create(buf, _i32(world), _i32(rank), C.byref(h)))
        obj = cls(h)
        obj.world, obj.rank = world, rank
        return obj

    @classmethod
    def from_torch_distributed(cls, group=None):
        """Bootstrap over an initialised torch.distributed group (plumbing only: rank 0's id is broadcast)."""
        import torch.distributed as dist
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        box = [cls.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0, group=group)
        return cls.create(box[0], world, rank)


SCORE_KEYS = ("sum", "points", "min", "max", "rays", "observations", "pairs_after_block_cull", "pairs")


def evaluate_trajectory_sharded(comm, views, max_distance, target_recon, width, height, bvh_tree, cloud, obs_rays,
                                recons, wrecons, h_recons=None, h_num_rows=None, stream=None):
    """uavmvs_evaluate_trajectory_sharded: this rank's shard + ONE collective for the trajectory score.
    `comm` may be None (single GPU).  Returns (score dict over all shards, (local sum, min, max))."""
    score, loc = (C.c_double * 8)(), (C.c_float * 3)()
    on_dev = isinstance(views, DeviceViews)
    if on_dev:
        vptr, nv = views.ptr, views.num_views
    else:
        views = _f(views, (-1, VIEW_FLOATS))
        vptr, nv = _p(views), views.shape[0]
    for a, dt in ((h_recons, np.float32), (h_num_rows, np.uint32)):
        assert a is None or (a.dtype == dt and a.flags.c_contiguous)
    _check(_lib.uavmvs_evaluate_trajectory_sharded(
        _stream_ptr(stream), None if comm is None else comm.ptr, vptr, _i32(1 if on_dev else 0), _u32(nv),
        _f32(max_distance), _f32(target_recon), _i32(width), _i32(height), bvh_tree.ptr, cloud.ptr, obs_rays.ptr,
        recons.ptr, wrecons.ptr, _p(h_recons), _p(h_num_rows), score, loc))
    out = dict(zip(SCORE_KEYS, (float(x) for x in score)))
    for k in ("points", "rays", "observations", "pairs_after_block_cull", "pairs"):
        out[k] = int(round(out[k]))
    out["score"] = out["sum"] / max(out["points"], 1)       # avg_wrecon, optimize_trajectory.cu:529
    return out, (float(loc[0]), float(loc[1]), float(loc[2]))


def allreduce_array(comm, array, count=None, stream=None):
    """SUM all-reduce of the first `count` floats of a device Array (partial sphere histograms of a sharded cloud)."""
    _check(_lib.uavmvs_allreduce_array(_stream_ptr(stream), None if comm is None else comm.ptr, array.ptr,
                                       _u64(array.n if count is None else count)))


def allgather_array(comm, local, count, out, stream=None):
    """All-gather `count` floats per rank (shard order) into the device Array `out`."""
    _check(_lib.uavmvs_allgather_array(_stream_ptr(stream), None if comm is None else comm.ptr, local.ptr,
                                       _u64(count), out.ptr))


def launch_count(reset=True):
    """Kernel launches issued by the evaluation path in this process since the last reset."""
    n = _u64(0)
    _check(_lib.uavmvs_launch_count(C.byref(n), _i32(1 if reset else 0)))
    return n.value


def profile_k1_enable(on=True):
    _check(_lib.uavmvs_profile_k1_enable(_i32(1 if on else 0)))


def profile_k1():
    """Device ms of k_view_masks, caps + scan, k_cull_emit, k_trace, k_retire of this thread's last K1 launch."""
    ms = (C.c_float * 5)()
    _check(_lib.uavmvs_profile_k1(ms))
    return [float(x) for x in ms]


def profile_last_collective():
    """Device ms of the score collective (pack, all-gather, fold) of this thread's last sharded evaluation."""
    ms = _f32(0)
    _check(_lib.uavmvs_profile_last_collective(C.byref(ms)))
    return ms.value


def evaluate_trajectory(views, max_distance, target_recon, width, height, bvh_tree, cloud, obs_rays,
                        recons, wrecons, stream=None, want_stats=True):
    """apps/evaluate_trajectory/evaluate_trajectory.cu:125-202 as one call.
    `views` is a host array [V,28] (copied inside the call) or a DeviceViews.
    Returns (sum_wrecon, min, max, stats[4])."""
    out = (C.c_float * 3)()
    stats = (C.c_uint64 * 4)() if want_stats else None
    if isinstance(views, DeviceViews):
        _check(_lib.uavmvs_evaluate_trajectory_device(
            _stream_ptr(stream), views.ptr, _u32(views.num_views), _f32(max_distance), _f32(target_recon),
            _i32(width), _i32(height), bvh_tree.ptr, cloud.ptr, obs_rays.ptr, recons.ptr, wrecons.ptr, out, stats))
    else:
        views = _f(views, (-1, VIEW_FLOATS))
        _check(_lib.uavmvs_evaluate_trajectory(
            _stream_ptr(stream), _p(views), _u32(views.shape[0]), _f32(max_distance), _f32(target_recon),
            _i32(width), _i32(height), bvh_tree.ptr, cloud.ptr, obs_rays.ptr, recons.ptr, wrecons.ptr, out, stats))
    st = [int(x) for x in stats] if want_stats else None
    return float(out[0]), float(out[1]), float(out[2]), st


def evaluate_trajectory_to_host(views, max_distance, target_recon, width, height, bvh_tree, cloud, obs_rays,
                                recons, wrecons, h_recons=None, h_num_rows=None, stream=None, want_stats=True):
    """evaluate_trajectory + the per-point downloads that follow it in the app (evaluate_trajectory.cu:188-194),
    overlapped with the evaluation.  h_recons float32[N] / h_num_rows uint32[N]: host arrays (pinned for overlap)."""
    out = (C.c_float * 3)()
    stats = (C.c_uint64 * 4)() if want_stats else None
    views = _f(views, (-1, VIEW_FLOATS))
    n = cloud.num_points if hasattr(cloud, "num_points") else None
    for a, dt in ((h_recons, np.float32), (h_num_rows, np.uint32)):
        assert a is None or (a.dtype == dt and a.flags.c_contiguous and (n is None or a.size == n))
    _check(_lib.uavmvs_evaluate_trajectory_to_host(
        _stream_ptr(stream), _p(views), _u32(views.shape[0]), _f32(max_distance), _f32(target_recon),
        _i32(width), _i32(height), bvh_tree.ptr, cloud.ptr, obs_rays.ptr, recons.ptr, wrecons.ptr,
        _p(h_recons), _p(h_num_rows), out, stats))
    st = [int(x) for x in stats] if want_stats else None
    return float(out[0]), float(out[1]), float(out[2]), st


def profile_last():
    """Device ms of K1, K2, K3+K8, K9 of this thread's last evaluate_trajectory call."""
    ms = (C.c_float * 4)()
    _check(_lib.uavmvs_profile_last(ms))
    return [float(x) for x in ms]


def k1_counters(reset=True):
    """[shaft packets tested, packets certified, rays certified, rays traced] since the last reset."""
    out = (C.c_uint64 * 4)()
    _check(_lib.uavmvs_k1_counters(out, _i32(1 if reset else 0)))
    return [int(x) for x in out]


def populate_spherical_histogram(view_pos, max_distance, bvh_tree, cloud, kd_tree, sphere_hist, stream=None,
                                 target_recon=None, obs_rays=None, recons=None):
    """Both overloads of libs/eval/kernels.h:46-67: the 9-argument form is selected by passing
    target_recon, obs_rays and recons."""
    vp = _f(view_pos, 3)
    if obs_rays is None:
        _check(_lib.uavmvs_populate_spherical_histogram(_stream_ptr(stream), _p(vp), _f32(max_distance),
                                                        bvh_tree.ptr, cloud.ptr, kd_tree.ptr, sphere_hist.ptr))
    else:
        _check(_lib.uavmvs_populate_spherical_histogram_recon(
            _stream_ptr(stream), _p(vp), _f32(max_distance), _f32(target_recon), bvh_tree.ptr, cloud.ptr,
            kd_tree.ptr, obs_rays.ptr, recons.ptr, sphere_hist.ptr))


def evaluate_spherical_histogram(calib, width, height, kd_tree, sphere_hist, hist, stream=None):
    """libs/eval/kernels.h:69-79"""
    K = _f(calib, 9)
    _check(_lib.uavmvs_evaluate_spherical_histogram(_stream_ptr(stream), _p(K), _i32(width), _i32(height),
                                                    kd_tree.ptr, sphere_hist.ptr, hist.ptr))


def populate_spherical_histogram_batch(positions, max_distance, bvh_tree, cloud, kd_tree, sphere_hists, stream=None,
                                       target_recon=0.0, obs_rays=None, recons=None):
    """B candidate positions in one call; sphere_hists is an Array of >= B * bins floats (row per position).
    K4 when obs_rays/recons are None, K5 otherwise.  Contributions are added (null the array first)."""
    pos = _f(positions).reshape(-1, 3)
    _check(_lib.uavmvs_populate_spherical_histogram_batch(
        _stream_ptr(stream), _p(pos), _u32(pos.shape[0]), _f32(max_distance), _f32(target_recon), bvh_tree.ptr,
        cloud.ptr, kd_tree.ptr, None if obs_rays is None else obs_rays.ptr, None if recons is None else recons.ptr,
        sphere_hists.ptr))


def evaluate_spherical_histogram_batch(calib, width, height, kd_tree, sphere_hists, num_positions, image_width=128,
                                       image_height=45, images=None, want_best=True, stream=None):
    """Batched K6 + device arg-min/arg-max.  Returns (min, argmin, max, argmax) arrays of length B
    (arg = y * image_width + x, first extremum in row-major order) or None when want_best is False."""
    K = _f(calib, 9)
    B = int(num_positions)
    if want_best:
        vmin, vmax = np.empty(B, np.float32), np.empty(B, np.float32)
        imin, imax = np.empty(B, np.uint32), np.empty(B, np.uint32)
    else:
        vmin = vmax = imin = imax = None
    _check(_lib.uavmvs_evaluate_spherical_histogram_batch(
        _stream_ptr(stream), _p(K), _i32(width), _i32(height), kd_tree.ptr, sphere_hists.ptr, _u32(B),
        _u32(image_width), _u32(image_height), None if images is None else images.ptr,
        _p(vmin), _p(imin), _p(vmax), _p(imax)))
    return (vmin, imin, vmax, imax) if want_best else None


class NelderMead(_Handle):
    """Simplices of the optimiser's views on the device and one lock-step simplex-downhill step for a batch of them
    (uavmvs_nm_*: libs/opti/nelder_mead.h:29-121 with the objective of optimize_trajectory.cu:366-438)."""
    _destroy = "uavmvs_nm_destroy"

    @classmethod
    def create(cls, num_views, num_bins):
        h = _vp()
        _check(_lib.uavmvs_nm_create(_u32(num_views), _u32(num_bins), C.byref(h)))
        obj = cls(h)
        obj.num_views = num_views
        return obj

    def upload(self, simplices):
        s = _f(simplices, (self.num_views, 4, 3))
        _check(_lib.uavmvs_nm_upload(self.ptr, _p(s)))

    def step(self, active, min_distance, max_distance, target_recon, calib, width, height, bvh_tree, cloud, kd_tree, obs_rays,
             recons, airspace=None, stream=None):
        a = _u(active)
        _check(_lib.uavmvs_nm_step(_stream_ptr(stream), self.ptr, _p(a), _u32(a.shape[0]), _f32(min_distance), _f32(max_distance),
                                   _f32(target_recon), _p(_f(calib, 9)), _i32(width), _i32(height), bvh_tree.ptr,
                                   (airspace or bvh_tree).ptr, cloud.ptr, kd_tree.ptr, obs_rays.ptr, recons.ptr))

    def apply_views(self, device_views, stream=None):
        """Writes the poses of the last step's views into a DeviceViews holding them (in the step's order)."""
        _check(_lib.uavmvs_nm_apply_views(_stream_ptr(stream), self.ptr, device_views._d, _u32(device_views.num_views)))

    def download(self, stream=None):
        """-> dict(simplices [V,4,3], vid [V], value [V], vmin [V], vtheta [V], vphi [V]) (synchronises)."""
        n = self.num_views
        sx = np.empty((n, 4, 3), np.float32)
        vid = np.empty(n, np.uint32)
        val, vmin, vth, vph = (np.empty(n, np.float32) for _ in range(4))
        _check(_lib.uavmvs_nm_download(_stream_ptr(stream), self.ptr, _p(sx), _p(vid), _p(val), _p(vmin), _p(vth), _p(vph)))
        return dict(simplices=sx, vid=vid, value=val, vmin=vmin, vtheta=vth, vphi=vph)


def estimate_capture_difficulty(max_distance, bvh_tree, mesh_size, kd_tree, cloud, stream=None):
    """libs/eval/kernels.h:81-84"""
    _check(_lib.uavmvs_estimate_capture_difficulty(_stream_ptr(stream), _f32(max_distance), bvh_tree.ptr,
                                                   _u32(mesh_size), kd_tree.ptr, cloud.ptr))


def sphere_vertices(radius=1.0, subdivisions=3):
    """generate_sphere_mesh(radius, subdivisions)'s vertex list (libs/geom/sphere.h:17-115): [1272, 3] at 3 subdivisions."""
    n = _u32(0)
    _check(_lib.uavmvs_sphere_vertices(_f32(radius), _u32(subdivisions), None, _u32(0), C.byref(n)))
    out = np.empty((n.value, 3), np.float32)
    _check(_lib.uavmvs_sphere_vertices(_f32(radius), _u32(subdivisions), _p(out), _u32(n.value), C.byref(n)))
    return out


def sample_proxy_cloud(verts, faces, samples_per_unit_area):
    """generate_proxy_cloud.cpp:84-184 on the device -> (points [n,3], normals [n,3]) float32, Z-order sorted."""
    verts, faces = _f(verts, (-1, 3)), _u(faces).reshape(-1, 3)
    h, n = _vp(), _u32(0)
    _check(_lib.uavmvs_proxy_cloud_sample(_p(verts), _u32(verts.shape[0]), _p(faces), _u32(faces.shape[0]),
                                          _f32(samples_per_unit_area), C.byref(h), C.byref(n)))
    pts, nrm = np.empty((n.value, 3), np.float32), np.empty((n.value, 3), np.float32)
    try:
        _check(_lib.uavmvs_proxy_cloud_download(h, _p(pts), _p(nrm)))
    finally:
        _lib.uavmvs_proxy_cloud_destroy(h)
    return pts, nrm
