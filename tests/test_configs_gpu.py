"""BASELINE.json configs 3, 4 and 5 at CONFIG SIZE on the GPU against the oracle (the toy-scene versions live in
test_hist_gpu.py / test_parity_gpu.py).

config 4  generate_guidance_volume.cu:222-277 on the config-2 scene (1M triangles, 1M points): 64 candidate voxels
          sampled from the free space of the 128^3 grid through the batched K4 + K6; per bin gate = sum over its
          contributions of 1e-5 |score| + 1e-6, K6 images, arg-min / arg-max.
config 3  optimize_trajectory.cu:313-428 on the resident config-2 state: remove 16 views, K2 + K3, 64 candidate
          positions through the batched K5 + K6 + device arg-min.
config 5  10M-triangle city, 8M points: the GPU-built tree is exported and walked by the oracle over a 64-view
          sample x all points: rays, un-clamped counters and stored rows bit-exact.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2(oracle):
    from uavmvs_b200 import api, apps, scenes
    assert api.device_count() >= 1
    sc = scenes.urban(1_000_000, 1_000_000, 2000, 256.0, seed=1)
    bvh = api.BVHTree.build_sah(sc.verts, sc.faces)                      # the tree bench.py times
    nodes, aabbs, tris, indices = bvh.export()
    obvh = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    sv = api.sphere_vertices(1.0, 3)
    assert np.array_equal(sv.view(np.uint32), oracle.sphere_vertices(1.0, 3).view(np.uint32))
    okd = oracle.KD.build(sv)
    kn, kv = okd.export()
    gkd = api.KDTree.create(kn, kv)                                      # same KD numbering on both sides: same bins
    rng = np.random.default_rng(4)
    values = rng.uniform(0.0, 1.2, sc.num_points).astype(np.float32)
    quals = rng.uniform(0.0, 1.0, sc.num_points).astype(np.float32)
    return dict(api=api, apps=apps, sc=sc, bvh=bvh, obvh=obvh, sv=sv, okd=okd, gkd=gkd, values=values, quals=quals)


def _bin_gate(bins, contrib, S):
    tol = np.zeros(S, np.float64)
    ok = bins != 0xFFFFFFFF
    np.add.at(tol, bins[ok], 1e-5 * np.abs(contrib[ok].astype(np.float64)) + 1e-6)
    return tol + 1e-12


def _check_images(api, oracle, c, hists_gpu, B):
    """Batched K6 + device extrema of B histogram rows against the oracle's K6 on the same rows."""
    sc, S = c["sc"], c["sv"].shape[0]
    rows = hists_gpu.download()[:B * S].reshape(B, S)
    images = api.Array.create(B * 45 * 128)
    vmin, imin, vmax, imax = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, c["gkd"], hists_gpu, B,
                                                                    images=images)
    img = images.download().reshape(B, 45 * 128)
    assert np.array_equal(vmin, img.min(1)) and np.array_equal(vmax, img.max(1))
    assert np.array_equal(imin, img.argmin(1).astype(np.uint32)) and np.array_equal(imax, img.argmax(1).astype(np.uint32))
    same_argmin = same_argmax = 0
    for b in range(B):
        i_ref = oracle.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, c["sv"], rows[b]).reshape(-1)
        tol = 1e-5 * np.abs(rows[b]).sum() + 1e-12
        assert np.all(np.abs(img[b].astype(np.float64) - i_ref) <= tol), b
        # the extremum's INDEX can only differ where the oracle's image has another pixel within the tolerance of its best
        for got, ref_idx in ((imin[b], int(i_ref.argmin())), (imax[b], int(i_ref.argmax()))):
            if got != ref_idx:
                assert abs(float(i_ref[got]) - float(i_ref[ref_idx])) <= 2 * tol, (b, got, ref_idx)
        same_argmin += int(imin[b] == int(i_ref.argmin()))
        same_argmax += int(imax[b] == int(i_ref.argmax()))
    return same_argmin, same_argmax


def test_config4_guidance_volume_voxels_at_config_size(c2, oracle):
    c, api, apps, sc = c2, c2["api"], c2["apps"], c2["sc"]
    S = c["sv"].shape[0]
    cloud = api.PointCloud.create(sc.points, sc.normals, c["values"], c["quals"])
    vol = apps.GuidanceVolume(sc, c["bvh"], cloud, c["gkd"], S, grid=128, batch=64)
    assert vol.free_total > 1_500_000
    B = vol.subsample(64)
    pos = vol.positions
    vol.hists.null()
    api.populate_spherical_histogram_batch(pos, sc.max_distance, c["bvh"], cloud, c["gkd"], vol.hists)
    h = vol.hists.download().reshape(B, S)
    nonzero = 0
    for b in range(B):
        h_ref, bins, scores = oracle.populate_spherical_histogram(c["obvh"], c["okd"], pos[b], sc.max_distance, sc.points,
                                                                  sc.normals, c["values"], c["quals"], S)
        assert np.all(np.abs(h[b].astype(np.float64) - h_ref) <= _bin_gate(bins, scores, S)), b
        nonzero += int(h_ref.any())
    assert nonzero >= 16                                     # the sample reaches positions that see the scene
    same_min, same_max = _check_images(api, oracle, c, vol.hists, B)
    assert same_max >= B - 2
    # a second run gives the same bits (order-independent fixed-point retire)
    vol.hists.null()
    api.populate_spherical_histogram_batch(pos, sc.max_distance, c["bvh"], cloud, c["gkd"], vol.hists)
    assert np.array_equal(vol.hists.download().view(np.uint32).reshape(B, S), h.view(np.uint32))


def test_config3_optimizer_candidates_at_config_size(c2, oracle):
    c, api, apps, sc = c2, c2["api"], c2["apps"], c2["sc"]
    S = c["sv"].shape[0]
    cloud = api.PointCloud.create(sc.points, sc.normals)
    loop = apps.OptimizerLoop(sc, c["bvh"], cloud, c["gkd"], S, want_views=16, candidates_per_view=4)
    loop.populate()
    loop.relax_independence()
    sel = loop.select()
    assert len(sel) == 16
    views = loop.views
    # oracle state: all views, minus the selected ones, K2
    obs_o = oracle.ObsRays(sc.num_points, 32)
    oracle.update_observation_rays(c["obvh"], True, views, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    oracle.update_observation_rays(c["obvh"], False, views[sel], sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    api.update_observation_rays_batch(False, views[sel], sc.max_distance, sc.width, sc.height, c["bvh"], cloud, loop.obs)
    loop.reeval()
    d_g, rows_g = loop.obs.download()
    assert np.array_equal(rows_g, obs_o.num_rows)                                        # counts after the removal: exact
    m = np.arange(32)[:, None] < np.minimum(rows_g, 32)[None, :]
    # rows after K2: the sorted keys are exact; the rows themselves are exact except among EQUAL keys -- the reference's
    # network is not idempotent on ties (kernels.cu:241 never exchanges equal keys, so a second K2 over an unchanged
    # column may permute its tied rows) and the oracle, like the reference app, re-sorts every column while the
    # incremental re-evaluation only visits the 32-point blocks K1 touched
    assert np.array_equal(d_g[:, :, 2].view(np.uint32)[m], obs_o.data[:, :, 2].view(np.uint32)[m])
    differ = (d_g.view(np.uint32) != obs_o.data.view(np.uint32)).any(axis=2) & m
    assert differ.mean() < 1e-4

    def colkeys(d):
        a = np.where(m[:, :, None], d, 0).view(np.uint32).astype(np.uint64)
        return np.sort(a[:, :, 0] * 7 + a[:, :, 1] * 13 + a[:, :, 2] * 31 + a[:, :, 3], axis=0)
    assert np.array_equal(colkeys(d_g), colkeys(obs_o.data))                            # same multiset of rows per column
    r_g = loop.recons.download()
    r_o = oracle.evaluate_observation_rays(obs_o)
    assert np.all(np.abs(r_g.astype(np.float64) - r_o) <= 1e-5 * np.maximum(np.abs(r_g), np.abs(r_o)) + 1e-6)
    rng = np.random.default_rng(11)
    cand = (views[sel, None, :3] + rng.normal(0.0, 2.0, (16, 4, 3))).reshape(-1, 3).astype(np.float32)
    B = cand.shape[0]
    hists = api.Array.create(B * S)
    hists.null()
    api.populate_spherical_histogram_batch(cand, sc.max_distance, c["bvh"], cloud, c["gkd"], hists,
                                           target_recon=sc.target_recon, obs_rays=loop.obs, recons=loop.recons)
    h = hists.download().reshape(B, S)
    for b in range(B):
        # the oracle consumes the GPU's recons so that the only differences are this kernel's
        h_ref, bins, deltas = oracle.populate_spherical_histogram_recon(c["obvh"], c["okd"], cand[b], sc.max_distance,
                                                                        sc.target_recon, sc.points, sc.normals, obs_o, r_g, S)
        assert np.all(np.abs(h[b].astype(np.float64) - h_ref) <= _bin_gate(bins, deltas, S)), b
        assert (h_ref < 0).any()
    same_min, _ = _check_images(api, oracle, c, hists, B)
    assert same_min >= B - 2


def test_config5_city_tree_counts_and_rows_bit_exact(oracle):
    from uavmvs_b200 import api, scenes
    sc = scenes.urban(10_000_000, 8_000_000, 10_000, 800.0, seed=1)
    bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
    nodes, aabbs, tris, indices = bvh.export()
    assert indices.shape[0] >= 10_000_000
    obvh = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    del nodes, aabbs, tris, indices
    vidx = np.unique(np.linspace(0, sc.num_views - 1, 64).round().astype(int))
    views = np.ascontiguousarray(sc.views28()[vidx])
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    rc, wr = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
    s_sum, _, _, stats = api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height, bvh, cloud, obs, rc, wr)
    obs_o = oracle.ObsRays(sc.num_points, 32)
    _, cnt = oracle.update_observation_rays(obvh, True, views, sc.max_distance, sc.width, sc.height, sc.points, sc.normals,
                                            obs_o, want_counters=True)
    oracle.process_observation_rays(obs_o)
    assert stats[0] == int(cnt[0]) > 100_000
    data, rows = obs.download()
    assert np.array_equal(rows, obs_o.num_rows) and int(rows.sum()) == stats[1]
    m = np.arange(32)[:, None] < np.minimum(rows, 32)[None, :]
    assert np.array_equal(data.view(np.uint32)[m], obs_o.data.view(np.uint32)[m])
    r_o = oracle.evaluate_observation_rays(obs_o)
    r_g = rc.download()
    # 64 of 10 000 views: most points hold two or three observations, where 1 - logistic cancels (SURVEY.md 7E gate)
    assert np.all(np.abs(r_g.astype(np.float64) - r_o) <= 1e-5 * np.maximum(np.abs(r_g), np.abs(r_o)) + 1e-6)
