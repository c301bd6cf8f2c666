"""File formats either side of the path (uavmvs_b200/formats.py): round trips, the header arithmetic of
libs/util/numpy_io.h, the bundled reference mesh, and a file-driven evaluation equal to the array-driven one."""
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
