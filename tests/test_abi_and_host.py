"""CPU-side checks: the C-ABI library loads and exports every symbol include/uavmvs_b200.h
declares (no compute calls without a GPU), the host mirror fails loudly without a device,
scene generators are deterministic, and the N>1 sharding logic under gloo (world size 2)."""
import ctypes
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from uavmvs_b200 import build
    return build.build()


def test_library_exports_every_declared_symbol(built):
    from uavmvs_b200 import abi
    names = abi.declared_symbols()
    assert len(names) >= 40 and "uavmvs_update_observation_rays" in names and "uavmvs_evaluate_trajectory" in names
    assert abi.missing_symbols() == []


def test_header_compiles_as_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "uavmvs_b200.h"\nint main(void) { return UAVMVS_VIEW_FLOATS == 28 ? 0 : 1; }\n')
    import subprocess
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                           "-o", str(tmp_path / "t.o")])


def test_no_cpu_fallback(built):
    """Without a device every entry point reports an error instead of computing on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from uavmvs_b200 import api
    with pytest.raises(api.UavmvsError):
        api.PointCloud.create(np.zeros((4, 3), np.float32), np.zeros((4, 3), np.float32))
    with pytest.raises(api.UavmvsError):
        api.VectorArray.create(4, 32)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "uavmvs_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "import oracle" not in text and "liboracle" not in text and "libref_acc" not in text, fn


def test_scenes_are_deterministic():
    from uavmvs_b200 import scenes
    a, b = scenes.tiny(), scenes.tiny()
    for name in ("verts", "faces", "points", "normals", "view_pos", "w2c", "calib"):
        assert np.array_equal(getattr(a, name), getattr(b, name))
    assert a.views28().shape == (a.num_views, 28) and a.views28().dtype == np.float32
    assert np.allclose(np.linalg.norm(a.normals, axis=1), 1.0, atol=1e-5)
    c = scenes.config1(num_points=2000)
    assert c.num_tris == 20482 and 150 <= c.num_views <= 260


def test_shard_ranges_cover_the_cloud():
    from uavmvs_b200.sharded import shard_range
    for n in (1, 7, 1000, 1_000_003):
        for world in (1, 2, 3, 8):
            r = [shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            assert max(hi - lo for lo, hi in r) - min(hi - lo for lo, hi in r) <= (n + world - 1) // world


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import oracle
    from uavmvs_b200 import scenes, sharded
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    sc = scenes.tiny()
    obvh = oracle.BVH.build(sc.verts, sc.faces, threads=1)      # replicated mesh
    lo, hi = sharded.shard_range(sc.num_points, rank, world)
    obs = oracle.ObsRays(hi - lo, 32)                           # the oracle stands in for the GPU shard here
    _, cnt = oracle.update_observation_rays(obvh, True, sc.views28(), sc.max_distance, sc.width, sc.height,
                                            sc.points[lo:hi], sc.normals[lo:hi], obs, want_counters=True)
    oracle.process_observation_rays(obs)
    recons = oracle.evaluate_observation_rays(obs)
    w = oracle.calculate_func_recons(recons, sc.target_recon)
    red = oracle.reduce(w)
    res = sharded.combine_scores(red[0], red[1], red[2], hi - lo, [int(cnt[0]), int(obs.num_rows.sum()), 0, 0])
    full = sharded.gather_point_vector(recons, sc.num_points)
    counts = sharded.gather_point_vector(obs.num_rows, sc.num_points)
    import torch
    h = torch.full((2, 8), float(rank + 1), dtype=torch.float32)
    sharded.allreduce_histograms(h)
    if rank == 0:
        np.savez(os.path.join(out_dir, "r.npz"), score=res["score"], mn=res["min"], mx=res["max"], pts=res["points"],
                 rays=res["stats"][0], obs=res["stats"][1], full=full, counts=counts, h=h.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("chunk", [1024, 8192, 0])
@pytest.mark.parametrize("world", [2, 8])
def test_bench_strong_split_is_a_partition(chunk, world):
    """bench.py's strong-scaling split (--shard-chunk: interleaved chunks with a rotating owner, 0 = contiguous Z-order
    ranges): every point of the cloud belongs to exactly one rank, ragged tail included, and the ranks are balanced."""
    import types
    import bench
    n = 100_000 + 37
    pts = np.arange(3 * n, dtype=np.float32).reshape(n, 3)                    # a point's x identifies it
    sc = types.SimpleNamespace(points=pts, normals=-pts, num_points=n)
    a = types.SimpleNamespace(scaling="strong", shard_chunk=chunk)
    seen = np.zeros(n, np.int32)
    sizes = []
    for r in range(world):
        p, nr = bench.rank_cloud(a, sc, r, world)
        assert p.shape == nr.shape and np.array_equal(nr, -p) and p.flags.c_contiguous
        ids = (p[:, 0] / 3).astype(np.int64)
        assert np.array_equal(pts[ids], p)
        if chunk == 0:
            assert np.all(np.diff(ids) == 1)                                   # one contiguous range
        seen[ids] += 1
        sizes.append(len(ids))
    assert np.all(seen == 1)
    assert max(sizes) - min(sizes) <= max(chunk, 32) * 2
    # one rank: the whole cloud, untouched
    p, _ = bench.rank_cloud(a, sc, 0, 1)
    assert p is pts


def test_point_sharding_world2_gloo(tmp_path, oracle, tiny_scene, tiny_tree):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_gloo_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r = np.load(tmp_path / "r.npz")
    sc = tiny_scene
    obvh, _ = tiny_tree
    obs = oracle.ObsRays(sc.num_points, 32)
    _, cnt = oracle.update_observation_rays(obvh, True, sc.views28(), sc.max_distance, sc.width, sc.height,
                                            sc.points, sc.normals, obs, want_counters=True)
    oracle.process_observation_rays(obs)
    recons = oracle.evaluate_observation_rays(obs)
    w = oracle.calculate_func_recons(recons, sc.target_recon)
    assert int(r["pts"]) == sc.num_points and int(r["rays"]) == int(cnt[0]) and int(r["obs"]) == int(obs.num_rows.sum())
    assert np.array_equal(r["full"].view(np.uint32), recons.view(np.uint32))     # shards are independent
    assert np.array_equal(r["counts"], obs.num_rows)
    assert abs(float(r["score"]) - float(w.astype(np.float64).mean())) < 1e-9
    assert float(r["mn"]) == float(w.min()) and float(r["mx"]) == float(w.max())
    assert np.all(r["h"] == 3.0)


def test_user_trees_are_validated_before_upload():
    """uavmvs_bvh_create_from_arrays / uavmvs_kdtree_create check user topology on the host (indices in range, a tree)
    and fail with a status code and a message -- before touching the device, so this runs without a GPU."""
    from uavmvs_b200 import api
    NAI = 0xFFFFFFFF
    aabbs = np.zeros((3, 6), np.float32)
    tris = np.zeros((2, 9), np.float32)
    idx = np.arange(2, dtype=np.uint32)
    good = np.array([[0, 2, 1, 2], [0, 1, NAI, NAI], [1, 2, NAI, NAI]], np.uint32)
    for bad, what in ((np.array([[0, 2, 1, 7], [0, 1, NAI, NAI], [1, 2, NAI, NAI]], np.uint32), "child index"),
                      (np.array([[0, 2, 1, 1], [0, 1, NAI, NAI], [1, 2, NAI, NAI]], np.uint32), "reachable twice"),
                      (np.array([[0, 2, 1, 0], [0, 1, NAI, NAI], [1, 2, NAI, NAI]], np.uint32), "reachable twice"),
                      (np.array([[0, 2, 1, 2], [0, 5, NAI, NAI], [1, 2, NAI, NAI]], np.uint32), "triangle range")):
        with pytest.raises(api.UavmvsError, match=what):
            api.BVHTree.create(bad, aabbs, tris, idx)
    verts = np.zeros((3, 3), np.float32)
    for bad, what in ((np.array([[0, 0, 1, 2], [1, 1, NAI, NAI], [9, 2, NAI, NAI]], np.uint32), "vertex outside"),
                      (np.array([[0, 3, 1, 2], [1, 1, NAI, NAI], [2, 2, NAI, NAI]], np.uint32), "dimension"),
                      (np.array([[0, 0, 1, 5], [1, 1, NAI, NAI], [2, 2, NAI, NAI]], np.uint32), "child index"),
                      (np.array([[0, 0, 1, 1], [1, 1, NAI, NAI], [2, 2, NAI, NAI]], np.uint32), "reachable twice")):
        with pytest.raises(api.UavmvsError, match=what):
            api.KDTree.create(bad, verts)
    # a chain deeper than the device search stack
    n = 64
    chain = np.array([[i, i % 3, i + 1 if i + 1 < n else NAI, NAI] for i in range(n)], np.uint32)
    with pytest.raises(api.UavmvsError, match="deeper than the search stack"):
        api.KDTree.create(chain, np.zeros((n, 3), np.float32))
    assert good.shape == (3, 4)


def test_sphere_vertices_equal_the_oracle(oracle):
    """uavmvs_sphere_vertices (host code of the product) against the oracle's restatement of libs/geom/sphere.h."""
    from uavmvs_b200 import api
    for sub, count in ((0, 12), (1, 72), (2, 312), (3, 1272)):
        v = api.sphere_vertices(1.0, sub)
        assert v.shape == (count, 3)
        assert np.array_equal(v.view(np.uint32), oracle.sphere_vertices(1.0, sub).view(np.uint32))
    v = api.sphere_vertices(1.0, 3)
    assert np.unique(v, axis=0).shape[0] == 642                     # 630 coincident twins
    assert np.array_equal(api.sphere_vertices(2.5, 2).view(np.uint32), oracle.sphere_vertices(2.5, 2).view(np.uint32))
