"""Pins the oracle's tree steps to the REFERENCE'S OWN CODE.

oracle/_ref/libref_acc.so is libs/acc (bvh_tree.h, kd_tree.h, primitives.h) compiled
unmodified from /root/reference; oracle/liboracle.so is our restatement.  Live comparison
runs where _ref was built (the build container); the committed tests/golden/ref_*.npz
(made by tests/golden/make_golden.py from the same _ref library) run everywhere.
"""
import hashlib
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from conftest import make_rays  # noqa: E402
import make_golden as mg  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def _need_ref(oracle):
    if oracle.ref() is None:
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")


# ------------------------------------------------------------------ golden (reference outputs)
def test_golden_sah_tree_and_rays(oracle, tiny_scene):
    g = np.load(os.path.join(GOLD, "ref_acc_bvh.npz"))
    ob = oracle.BVH.build(tiny_scene.verts, tiny_scene.faces, threads=2)
    a = ob.export()
    assert a.num_nodes == int(g["tiny_num_nodes"])
    for name in ("nodes", "aabbs", "tris", "indices"):          # libs/acc/bvh_tree.h:335-399, bit for bit
        assert np.array_equal(_sha(getattr(a, name)), g["tiny_%s_sha" % name]), name
    rays = make_rays(tiny_scene, 20_000, seed=11)
    h, f, t = ob.intersect(rays, any_hit=False)                  # libs/acc/bvh_tree.h:419-467
    assert np.array_equal(np.packbits(h), g["tiny_hit"])
    assert np.array_equal(np.where(h > 0, f, 0xFFFFFFFF).astype(np.uint32), g["tiny_face"])
    assert np.array_equal(np.where(h > 0, t, 0).astype(np.float32).view(np.uint32), g["tiny_t"].view(np.uint32))
    assert np.array_equal(ob.intersect(rays, any_hit=True)[0], h)


def test_golden_slack_zone_rays(oracle):
    """Rays within +-5 mm of a silhouette edge: which triangle (roof, wall, ground) wins
    depends on the 1e-3 barycentric slack and on the tree (libs/acc/primitives.h:239-242)."""
    g = np.load(os.path.join(GOLD, "ref_acc_bvh.npz"))
    mv, mf = mg.micro_mesh()
    ob = oracle.BVH.build(mv, mf, threads=1)
    a = ob.export()
    assert np.array_equal(a.nodes, g["micro_nodes"]) and np.array_equal(a.aabbs, g["micro_aabbs"])
    assert np.array_equal(a.tris, g["micro_tris"]) and np.array_equal(a.indices, g["micro_indices"])
    er = mg.edge_rays(4000, 5)
    h, f, _ = ob.intersect(er, any_hit=False)
    assert np.array_equal(np.packbits(h), g["micro_hit"])
    face = np.where(h > 0, f, 0xFFFFFFFF).astype(np.uint32)
    assert np.array_equal(face, g["micro_face"])
    assert len(np.unique(face)) >= 3                              # roof, walls and ground all occur


def test_golden_sphere_kdtree(oracle):
    g = np.load(os.path.join(GOLD, "ref_acc_kd.npz"))
    sv = oracle.sphere_vertices(1.0, 3)
    assert sv.shape == (1272, 3) and len(np.unique(sv, axis=0)) == 642    # libs/geom/sphere.h:60-79 quirk
    assert np.array_equal(_sha(sv), g["sphere_sha"])
    kd = oracle.KD.build(sv)
    nodes, verts = kd.export()
    assert np.array_equal(nodes, g["nodes"])                      # libs/acc/kd_tree.h:92-158, single thread
    assert np.array_equal(_sha(verts), g["verts_sha"])
    q = np.random.default_rng(3).normal(size=(5000, 3)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    idx, dist = kd.find_nn(q)
    assert np.array_equal(dist.view(np.uint32), g["nn_dist"].view(np.uint32))
    # coincident duplicate vertices: the winner among twins depends on the traversal; positions agree
    assert np.array_equal(sv[idx], sv[g["nn_idx"].astype(np.int64)])


# ------------------------------------------------------------------ live (reference code run here)
def test_live_sah_build_equals_reference(oracle):
    _need_ref(oracle)
    from uavmvs_b200 import scenes
    for sc in (scenes.tiny(), scenes.dense()):
        a = oracle.BVH.build(sc.verts, sc.faces, threads=2).export()
        r = oracle.RefBVH(sc.verts, sc.faces, threads=2).export()
        for name in ("nodes", "aabbs", "tris", "indices"):
            assert np.array_equal(getattr(a, name), getattr(r, name)), (sc.name, name)


def test_live_binned_sah_path_equals_reference(oracle):
    """> 64 primitives per node takes the binned-SAH branch (libs/acc/bvh_tree.h:160-175)."""
    _need_ref(oracle)
    from uavmvs_b200 import scenes
    verts, faces, _ = scenes.urban_mesh(60_000, 128.0, seed=21)
    a = oracle.BVH.build(verts, faces, threads=4).export()
    r = oracle.RefBVH(verts, faces, threads=4).export()
    for name in ("nodes", "aabbs", "tris", "indices"):
        assert np.array_equal(getattr(a, name), getattr(r, name)), name


def test_live_intersect_equals_reference(oracle):
    _need_ref(oracle)
    from uavmvs_b200 import scenes
    sc = scenes.dense()
    ob = oracle.BVH.build(sc.verts, sc.faces, threads=2)
    rb = oracle.RefBVH(sc.verts, sc.faces, threads=2)
    rays = make_rays(sc, 200_000, seed=4)
    h, f, t = ob.intersect(rays, any_hit=False)
    hr, fr, tr = rb.intersect(rays)
    assert np.array_equal(h, hr)
    assert np.array_equal(f[h > 0], fr[h > 0])
    assert np.array_equal(t[h > 0].view(np.uint32), tr[h > 0].view(np.uint32))
    assert np.array_equal(ob.intersect(rays, any_hit=True)[0], hr)  # any-hit early exit: same boolean


def test_live_degenerate_and_axis_parallel(oracle):
    """Flat axis (NaN bin index cast, bvh_tree.h:199,240), zero direction components."""
    _need_ref(oracle)
    rng = np.random.default_rng(8)
    n = 300
    xy = rng.uniform(0, 10, (n, 3, 2))
    verts = np.concatenate([xy, np.zeros((n, 3, 1))], 2).reshape(-1, 3).astype(np.float32)   # all z = 0
    faces = np.arange(3 * n, dtype=np.uint32).reshape(-1, 3)
    ob = oracle.BVH.build(verts, faces, threads=1)
    rb = oracle.RefBVH(verts, faces, threads=1)
    a, r = ob.export(), rb.export()
    assert np.array_equal(a.nodes, r.nodes) and np.array_equal(a.aabbs, r.aabbs)
    o = np.concatenate([rng.uniform(0, 10, (2000, 2)), np.full((2000, 1), 5.0)], 1)
    d = np.tile([0.0, 0.0, -1.0], (2000, 1))
    rays = np.concatenate([o, d, np.zeros((2000, 1)), np.full((2000, 1), 10.0)], 1).astype(np.float32)
    h, f, t = ob.intersect(rays, any_hit=False)
    hr, fr, tr = rb.intersect(rays)
    assert np.array_equal(h, hr) and np.array_equal(f[h > 0], fr[h > 0])
    assert 0.1 < h.mean() < 1.0


def test_live_kdtree_equals_reference(oracle):
    _need_ref(oracle)
    import ctypes  # noqa: F401
    R = oracle.ref()
    sv = oracle.sphere_vertices(1.0, 3)
    rk = R.ref_kd_build(np.ascontiguousarray(sv), sv.shape[0], 1)
    n = R.ref_kd_num_nodes(rk)
    kn, kv = np.empty((n, 4), np.uint32), np.empty((n, 3), np.float32)
    R.ref_kd_export(rk, kn, kv)
    nodes, verts = oracle.KD.build(sv).export()
    assert np.array_equal(nodes, kn) and np.array_equal(verts, kv)
    q = np.random.default_rng(12).normal(size=(20000, 3)).astype(np.float32)
    qi, qd = np.empty(20000, np.uint32), np.empty(20000, np.float32)
    R.ref_kd_find_nn(rk, q, 20000, qi, qd)
    idx, dist = oracle.KD.build(sv).find_nn(q)
    R.ref_kd_destroy(rk)
    assert np.array_equal(dist.view(np.uint32), qd.view(np.uint32))
    assert np.array_equal(sv[idx], sv[qi])


# ------------------------------------------------------------------ closest point (libs/acc/bvh_tree.h:492-548)
def test_golden_closest_point(oracle, tiny_scene):
    import make_golden_closest as mc
    g = np.load(os.path.join(GOLD, "ref_acc_closest.npz"))
    ob = oracle.BVH.build(tiny_scene.verts, tiny_scene.faces, threads=2)
    q = mc.queries(tiny_scene)
    for k, md in enumerate(mc.MAX_DISTS):
        f, i, c = ob.closest_point(q, md)
        assert np.array_equal(np.packbits(f), g["found%d" % k])
        assert np.array_equal(i, g["idx%d" % k])                                  # leaf-order triangle index
        assert np.array_equal(c.view(np.uint32), g["cp%d" % k].view(np.uint32))   # the point, bit for bit
    assert 0 < np.unpackbits(g["found2"]).sum() < np.unpackbits(g["found1"]).sum() < len(q)


def test_live_closest_point_equals_reference(oracle):
    _need_ref(oracle)
    from uavmvs_b200 import scenes
    sc = scenes.dense()
    ob = oracle.BVH.build(sc.verts, sc.faces, threads=2)
    rb = oracle.RefBVH(sc.verts, sc.faces, 2)
    rng = np.random.default_rng(3)
    q = np.concatenate([rng.uniform([-5, -5, -5], [70, 70, 90], (20000, 3)),
                        sc.points + rng.normal(0, 0.05, sc.points.shape)]).astype(np.float32)
    for md in (float("inf"), 2.0):
        a, b = ob.closest_point(q, md), rb.closest_point(q, md)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        assert np.array_equal(a[2].view(np.uint32), b[2].view(np.uint32))
