"""File formats either side of the path (uavmvs_b200/formats.py): round trips, the header arithmetic of
libs/util/numpy_io.h, the bundled reference mesh, and a file-driven evaluation equal to the array-driven one."""
import ctypes as C
import os

import numpy as np
import pytest

from uavmvs_b200 import formats as F
from uavmvs_b200 import scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REF_PLY = "/root/reference/res/meshes/sphere.ply"


def test_trajectory_round_trip_and_views(tmp_path):
    sc = scenes.tiny()
    rot = sc.w2c[:, :3, :3]
    flen = np.full(sc.num_views, scenes.FLEN, np.float32)
    p = str(tmp_path / "t.traj")
    F.save_trajectory(p, sc.view_pos, rot, flen)
    pos2, rot2, flen2 = F.load_trajectory(p)
    assert np.array_equal(pos2, sc.view_pos) and np.array_equal(rot2, rot) and np.array_equal(flen2, flen)
    v = F.views_from_trajectory(pos2, rot2, flen2, sc.width, sc.height)
    ref = sc.views28()
    assert np.array_equal(v[:, :3], ref[:, :3])                       # position
    assert np.allclose(v[:, 19:], ref[:, 19:], rtol=1e-6, atol=0)     # flen * width in fp32 (MVE) vs fp64-then-rounded
    assert np.allclose(v[:, 3:19], ref[:, 3:19], rtol=0, atol=2e-5)   # -R pos in fp32 vs fp64-then-rounded
    open(p, "w").write("3\n1 2 3\n")
    with pytest.raises(ValueError):
        F.load_trajectory(p)                                          # "Invalid trajectory file" (:66-69)


def test_ply_round_trips_and_cloud_defaults(tmp_path):
    sc = scenes.tiny()
    m = str(tmp_path / "mesh.ply")
    F.save_ply(m, sc.verts, sc.faces)
    d = F.load_ply(m)
    assert np.array_equal(d["vertices"], sc.verts) and np.array_equal(d["faces"], sc.faces) and d["normals"] is None
    c = str(tmp_path / "cloud.ply")
    F.save_ply(c, sc.points, None, sc.normals)
    pts, nrm, val, qual = F.load_point_cloud(c)                       # libs/util/cio.h:45-46 defaults
    assert np.array_equal(pts, sc.points) and np.array_equal(nrm, sc.normals)
    assert not val.any() and np.all(qual == 1.0)
    vals = np.linspace(0, 1, sc.num_points, dtype=np.float32)
    F.save_ply(c, sc.points, None, sc.normals, vals, 1.0 - vals)
    _, _, val, qual = F.load_point_cloud(c)
    assert np.array_equal(val, vals) and np.array_equal(qual, 1.0 - vals)
    with pytest.raises(ValueError):
        F.load_point_cloud(m)                                         # "Point cloud has no vertex normals."
    # ascii flavour
    a = str(tmp_path / "a.ply")
    with open(a, "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment x\nelement vertex 3\nproperty float x\nproperty float y\nproperty float z\n"
                "element face 1\nproperty list uchar int vertex_index\nend_header\n0 0 0\n1 0 0\n0 1 0.5\n3 0 1 2\n")
    d = F.load_ply(a)
    assert d["vertices"].shape == (3, 3) and d["faces"].tolist() == [[0, 1, 2]] and d["vertices"][2, 2] == 0.5


def test_numpy_file_is_what_the_reference_writes(tmp_path):
    p = str(tmp_path / "o.npy")
    h, e, o = np.arange(7, dtype=np.float32), np.ones(7, np.float32), np.full(7, 2.5, np.float32)
    F.save_numpy_file(p, h, e, o)                                     # evaluate_heuristic.cu:293
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x93NUMPY\x02\x00"
    hl = int.from_bytes(raw[8:12], "little")
    assert (12 + hl) % 16 == 0 and raw[12 + hl - 1:12 + hl] == b"\n"
    assert raw[12:12 + hl].startswith(b"{'descr': '<f4', 'fortran_order': False, 'shape': (3, 7), }")
    a = np.load(p)
    assert a.shape == (3, 7) and np.array_equal(a[0], h) and np.array_equal(a[2], o)
    F.save_numpy_file(p, np.arange(5, dtype=np.uint32))
    assert np.array_equal(np.load(p), np.arange(5, dtype=np.int32)[None])   # 'i4': the reference only knows f / i


def test_volume_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    imgs = {5: rng.random((45, 128), dtype=np.float32), 63: rng.random((45, 128), dtype=np.float32)}
    p = str(tmp_path / "g.vol")
    F.save_volume(p, 4, 4, 4, [0, -1, 2.5], [10, 11, 12], imgs)
    w, h, d, mn, mx, out = F.load_volume(p)
    assert (w, h, d) == (4, 4, 4) and mn.tolist() == [0, -1, 2.5] and mx.tolist() == [10, 11, 12]
    assert sorted(out) == [5, 63] and all(np.array_equal(out[k][:, :, 0], imgs[k]) for k in imgs)
    assert open(p, "rb").read().startswith(b"VOL 0.1\n4 4 4 \n")
    F.save_volume(p, 2, 2, 2, [0, 0, 0], [1, 1, 1], {})
    assert F.load_volume(p)[5] == {}


def test_bundled_sphere_fixture_is_the_reference_mesh():
    g = np.load(os.path.join(GOLD, "ref_sphere_mesh.npz"))
    assert g["vertices"].shape == (10242, 3) and g["faces"].shape == (20480, 3)
    r = np.linalg.norm(g["vertices"].astype(np.float64), axis=1)
    assert abs(r.min() - 1.0) < 1e-6 and abs(r.max() - 1.0) < 1e-6   # unit sphere
    sv, sf = scenes.bundled_sphere()
    assert np.array_equal(sv.astype(np.float32), g["vertices"]) and np.array_equal(sf, g["faces"])
    if os.path.exists(REF_PLY):                                       # build container only
        d = F.load_ply(REF_PLY)
        assert np.array_equal(d["vertices"], g["vertices"]) and np.array_equal(d["faces"], g["faces"])


def _geom_inputs():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden_geom", os.path.join(os.path.dirname(__file__), "golden", "make_golden_geom.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.inputs()


def test_writers_equal_the_reference_bytes(tmp_path):
    """.npy (libs/util/numpy_io.h:59-102) and VOL (libs/geom/volume_io.h:22-47) files byte for byte against what the
    reference's own writers produced (tests/golden/ref_geom.npz, made by oracle/_ref/libref_geom.so), and live against
    that library when it is present."""
    from uavmvs_b200 import formats
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_geom.npz"))
    f3, i1, (w, h, d, vmin, vmax, idx, imgs) = _geom_inputs()
    p = str(tmp_path / "a.npy")
    formats.save_numpy_file(p, f3[0], f3[1], f3[2])
    assert open(p, "rb").read() == g["npy_f32x3"].tobytes()
    assert np.array_equal(np.load(p), f3)                              # and numpy reads the reference's layout
    formats.save_numpy_file(p, i1[0])
    assert open(p, "rb").read() == g["npy_i32x1"].tobytes()
    pv = str(tmp_path / "a.vol")
    formats.save_volume(pv, w, h, d, vmin, vmax, {int(i): imgs[k] for k, i in enumerate(idx)})
    assert open(pv, "rb").read() == g["vol"].tobytes()
    W, H, D, mn, mx, images = formats.load_volume(pv)
    assert (W, H, D) == (w, h, d) and sorted(images) == [int(i) for i in idx]
    assert all(np.array_equal(images[int(i)], imgs[k]) for k, i in enumerate(idx))
    import oracle
    R = oracle.refgeom()
    if R is not None:                                                  # live: the reference's loader reads OUR file
        dims, rmn, rmx = np.zeros(3, np.uint32), np.zeros(3, np.float32), np.zeros(3, np.float32)
        n = C.c_uint32(0)
        ridx, rdata = np.zeros(8, np.uint32), np.zeros(imgs.size, np.float32)
        assert R.ref_load_volume(pv.encode(), dims, rmn, rmx, C.byref(n), ridx, 8, rdata, rdata.size) == 0
        assert list(dims) == [w, h, d] and n.value == len(idx) and np.array_equal(ridx[:n.value], idx)
        assert np.array_equal(rdata, imgs.reshape(-1))
        assert np.allclose(rmn, vmin, rtol=1e-5) and np.allclose(rmx, vmax, rtol=1e-5)   # six significant digits survive the text header


def test_sphere_vertices_equal_the_reference():
    """The 1272-vertex direction set: product (uavmvs_sphere_vertices) == oracle restatement == the reference's own
    generate_sphere_mesh(1.0f, 3u) (golden fixture; live when oracle/_ref/libref_geom.so is present)."""
    import oracle
    from uavmvs_b200 import api
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_geom.npz"))
    ref = g["sphere"]
    assert ref.shape == (1272, 3) and np.unique(ref, axis=0).shape[0] == 642
    assert np.array_equal(oracle.sphere_vertices(1.0, 3).view(np.uint32), ref.view(np.uint32))
    assert np.array_equal(api.sphere_vertices(1.0, 3).view(np.uint32), ref.view(np.uint32))
    R = oracle.refgeom()
    if R is not None:
        live = np.zeros((2048, 3), np.float32)
        n = R.ref_sphere_vertices(1.0, 3, live, 2048)
        assert n in (0, 1272)                                          # 0: generate_sphere_mesh only works once per process
        if n:
            assert np.array_equal(live[:n].view(np.uint32), ref.view(np.uint32))
