"""Spherical-histogram path (K4/K5/K6/K7) through the C ABI against the CPU oracle.

Bins (nearest sphere vertex through the same KD-tree) are integer work: bit-exact.  Histogram
sums are float accumulations in a different order (shared-memory + global atomics vs the oracle's
serial double sum): gated at 1e-5 relative to the magnitude of the summands, per bin.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from uavmvs_b200 import api as a
    assert a.device_count() >= 1
    return a


@pytest.fixture(scope="module")
def setup(api, oracle):
    from uavmvs_b200 import scenes
    sc = scenes.dense()
    obvh = oracle.BVH.build(sc.verts, sc.faces, threads=2)
    arrays = obvh.export()
    gbvh = api.BVHTree.create(arrays.nodes, arrays.aabbs, arrays.tris, arrays.indices)
    sv = oracle.sphere_vertices(1.0, 3)
    okd = oracle.KD.build(sv)
    kn, kv = okd.export()
    gkd = api.KDTree.create(kn, kv)
    rng = np.random.default_rng(9)
    values = rng.uniform(0.0, 1.2, sc.num_points).astype(np.float32)
    quals = rng.uniform(-0.1, 1.0, sc.num_points).astype(np.float32)
    cloud = api.PointCloud.create(sc.points, sc.normals, values, quals)
    return dict(sc=sc, obvh=obvh, gbvh=gbvh, sv=sv, okd=okd, gkd=gkd, values=values, quals=quals, cloud=cloud)


def _hist_close(h_gpu, h_ref, bins, contrib, S):
    mag = np.zeros(S, np.float64)
    ok = bins != 0xFFFFFFFF
    np.add.at(mag, bins[ok], np.abs(contrib[ok].astype(np.float64)))
    return np.all(np.abs(h_gpu.astype(np.float64) - h_ref) <= 1e-5 * mag + 1e-12)


def _k5_close(h_gpu, h_ref, bins, deltas, S):
    """K5's delta = func(new) - func(old) can cancel, so every contribution is gated at
    |a - b| <= 1e-5 max(|a|, |b|) + 1e-6; a bin is the sum of its contributions, hence of their gates."""
    tol = np.zeros(S, np.float64)
    ok = bins != 0xFFFFFFFF
    np.add.at(tol, bins[ok], 1e-5 * np.abs(deltas[ok].astype(np.float64)) + 1e-6)
    return np.all(np.abs(h_gpu.astype(np.float64) - h_ref) <= tol + 1e-12)


POSITIONS = [(32.0, 32.0, 45.0), (10.0, 50.0, 30.0), (60.0, 5.0, 70.0), (32.0, 32.0, 400.0)]


def test_populate_spherical_histogram_k4(api, oracle, setup):
    s = setup
    sc, S = s["sc"], s["sv"].shape[0]
    for pos in POSITIONS:
        h_ref, bins, scores = oracle.populate_spherical_histogram(s["obvh"], s["okd"], np.array(pos, np.float32),
                                                                  sc.max_distance, sc.points, sc.normals,
                                                                  s["values"], s["quals"], S)
        hist = api.Array.create(S)
        hist.null()
        api.populate_spherical_histogram(pos, sc.max_distance, s["gbvh"], s["cloud"], s["gkd"], hist)
        h_gpu = hist.download()
        assert np.array_equal(h_gpu != 0, h_ref != 0) or _hist_close(h_gpu, h_ref, bins, scores, S)
        assert _hist_close(h_gpu, h_ref, bins, scores, S)
        if pos[2] < 100:
            assert (bins != 0xFFFFFFFF).sum() > 100
        else:
            assert not h_gpu.any()                               # beyond max_distance: nothing populated


def test_populate_spherical_histogram_k5_and_image_k6(api, oracle, setup):
    s = setup
    sc, S = s["sc"], s["sv"].shape[0]
    # state after a full evaluation: obs rows sorted, recons
    obs_o = oracle.ObsRays(sc.num_points, 32)
    oracle.update_observation_rays(s["obvh"], True, sc.views28(), sc.max_distance, sc.width, sc.height,
                                   sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    r_o = oracle.evaluate_observation_rays(obs_o)
    obs = api.VectorArray.create(sc.num_points, 32)
    recons, wrecons = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
    api.evaluate_trajectory(sc.views28(), sc.max_distance, sc.target_recon, sc.width, sc.height, s["gbvh"],
                            s["cloud"], obs, recons, wrecons)
    assert np.array_equal(obs.download(data=False)[1], obs_o.num_rows)
    for pos in POSITIONS[:3]:
        # the oracle consumes the GPU's recons so that the only differences are this kernel's
        h_ref, bins, deltas = oracle.populate_spherical_histogram_recon(
            s["obvh"], s["okd"], np.array(pos, np.float32), sc.max_distance, sc.target_recon, sc.points, sc.normals,
            obs_o, recons.download(), S)
        hist = api.Array.create(S)
        hist.null()
        api.populate_spherical_histogram(pos, sc.max_distance, s["gbvh"], s["cloud"], s["gkd"], hist,
                                         target_recon=sc.target_recon, obs_rays=obs, recons=recons)
        h_gpu = hist.download()
        ok = bins != 0xFFFFFFFF
        assert _k5_close(h_gpu, h_ref, bins, deltas, S)
        assert ok.sum() > 100 and (h_ref < 0).any()
        # K6 on the GPU histogram vs the oracle on the same histogram
        img = api.Image.create(128, 45)
        api.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, s["gkd"], hist, img)
        i_gpu = img.download()
        i_ref = oracle.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, s["sv"], h_gpu)
        assert i_gpu.shape == (45, 128)
        scale = np.abs(h_gpu).sum()
        assert np.all(np.abs(i_gpu.astype(np.float64) - i_ref) <= 1e-5 * scale + 1e-12)
        # the optimiser's argmin (apps/optimize_trajectory/optimize_trajectory.cu:415-428) agrees in value
        assert abs(float(i_gpu.min()) - float(i_ref.min())) <= 1e-5 * scale + 1e-12


def test_estimate_capture_difficulty_k7(api, oracle, setup):
    s = setup
    sc = s["sc"]
    n = 512
    pts, nrm = sc.points[:n], sc.normals[:n]
    mesh_size = sc.num_tris // 2            # pretend the second half of the faces is "airspace"
    v_o, q_o = oracle.estimate_capture_difficulty(s["obvh"], sc.max_distance, mesh_size, s["sv"], pts, nrm)
    cloud = api.PointCloud.create(pts, nrm)
    api.estimate_capture_difficulty(sc.max_distance, s["gbvh"], mesh_size, s["gkd"], cloud)
    v_g, q_g = cloud.download_attributes()
    assert np.allclose(v_g, v_o, rtol=1e-5, atol=1e-6)
    assert np.allclose(q_g, q_o, rtol=1e-5, atol=1e-6)
    assert q_o.max() > 0 and v_o.min() < 1.5


def test_k7_wide_walk_equals_reference_walk_with_ties_and_airspace(api, oracle, monkeypatch):
    """K7 on the 4-wide tree (k7.cu) against the reference-order closest-hit walk (UAVMVS_K7_EXACT=1) and the oracle, on a
    tree built to need the fallback: the "airspace" (faces >= mesh_size) holds exact duplicates of every 5th mesh face
    (closest hits of the two classes at the SAME t: the visiting order decides) and a two-triangle sheet above the block
    whose diagonal is hit twice within the +-1e-3 barycentric slack; the points sit on axis-aligned faces, so a sixth of
    the icosphere directions have a zero component (axes dropped in the node test)."""
    from uavmvs_b200 import scenes
    sc = scenes.tiny()
    nv = sc.verts.shape[0]
    dup = sc.faces[::5]
    z = float(sc.verts[:, 2].max()) + 7.0
    lo, hi = sc.verts[:, :2].min(0) - 3.0, sc.verts[:, :2].max(0) + 3.0
    sheet_v = np.array([[lo[0], lo[1], z], [hi[0], lo[1], z], [hi[0], hi[1], z], [lo[0], hi[1], z]], np.float32)
    sheet_f = np.array([[nv, nv + 1, nv + 2], [nv, nv + 2, nv + 3]], sc.faces.dtype)
    verts = np.ascontiguousarray(np.concatenate([sc.verts, sheet_v]).astype(np.float32))
    faces = np.ascontiguousarray(np.concatenate([sc.faces, dup, sheet_f]).astype(sc.faces.dtype))
    mesh_size = sc.faces.shape[0]
    gbvh = api.BVHTree.build_sah(verts, faces)
    nodes, aabbs, tris, indices = gbvh.export()
    obvh = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    sv = api.sphere_vertices(1.0, 3)
    kn, kv = oracle.KD.build(sv).export()
    gkd = api.KDTree.create(kn, kv)
    n = 1536 + 7                                             # a ragged last block
    pts, nrm = np.ascontiguousarray(sc.points[:n]), np.ascontiguousarray(sc.normals[:n])
    out = {}
    for mode in ("wide", "exact", "wide_small_batches"):
        if mode == "exact":
            monkeypatch.setenv("UAVMVS_K7_EXACT", "1")
        else:
            monkeypatch.delenv("UAVMVS_K7_EXACT", raising=False)
        if mode == "wide_small_batches":
            monkeypatch.setenv("UAVMVS_K7_BATCH", "512")   # several launches, the last one ragged
        cloud = api.PointCloud.create(pts, nrm)
        api.estimate_capture_difficulty(sc.max_distance, gbvh, mesh_size, gkd, cloud)
        out[mode] = cloud.download_attributes()
    for mode in ("wide", "wide_small_batches"):
        assert np.array_equal(out[mode][0].view(np.uint32), out["exact"][0].view(np.uint32)), mode    # values: same bits
        assert np.array_equal(out[mode][1].view(np.uint32), out["exact"][1].view(np.uint32)), mode    # qualities: same bits
    v_o, q_o = oracle.estimate_capture_difficulty(obvh, sc.max_distance, mesh_size, sv, pts, nrm)
    assert np.allclose(out["wide"][0], v_o, rtol=1e-5, atol=1e-6) and np.allclose(out["wide"][1], q_o, rtol=1e-5, atol=1e-6)
    assert q_o.max() > 0
    # with mesh_size beyond every face the predicate is plain any-hit: the two walks agree there too
    monkeypatch.delenv("UAVMVS_K7_BATCH", raising=False)
    res = []
    for exact in (False, True):
        if exact:
            monkeypatch.setenv("UAVMVS_K7_EXACT", "1")
        else:
            monkeypatch.delenv("UAVMVS_K7_EXACT", raising=False)
        cloud = api.PointCloud.create(pts, nrm)
        api.estimate_capture_difficulty(sc.max_distance, gbvh, faces.shape[0], gkd, cloud)
        res.append(cloud.download_attributes())
    assert np.array_equal(res[0][0].view(np.uint32), res[1][0].view(np.uint32))
    assert np.array_equal(res[0][1].view(np.uint32), res[1][1].view(np.uint32))


def test_batched_k4_k6_equal_single_calls(api, oracle, setup):
    """uavmvs_populate_spherical_histogram_batch / _evaluate_spherical_histogram_batch (config 4: one call per
    batch of candidate voxels) against the per-position entry points and the oracle."""
    s = setup
    sc, S = s["sc"], s["sv"].shape[0]
    rng = np.random.default_rng(21)
    pos = np.concatenate([np.array(POSITIONS, np.float32),
                          rng.uniform([0, 0, 15], [64, 64, 90], (28, 3)).astype(np.float32)])
    B = pos.shape[0]
    hists = api.Array.create(B * S)
    hists.null()
    api.populate_spherical_histogram_batch(pos, sc.max_distance, s["gbvh"], s["cloud"], s["gkd"], hists)
    h_b = hists.download().reshape(B, S)
    for b in range(B):
        h_ref, bins, scores = oracle.populate_spherical_histogram(s["obvh"], s["okd"], pos[b], sc.max_distance,
                                                                  sc.points, sc.normals, s["values"], s["quals"], S)
        assert _hist_close(h_b[b], h_ref, bins, scores, S), b
        assert np.array_equal(h_b[b] != 0, h_ref != 0) or _hist_close(h_b[b], h_ref, bins, scores, S)
    assert not h_b[3].any() and (h_b != 0).sum() > 1000
    # batched K6: images bit-identical to the per-position kernel on the same histogram rows (same additions,
    # same order), extrema = first occurrence in row-major order
    images = api.Array.create(B * 45 * 128)
    vmin, imin, vmax, imax = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, s["gkd"], hists, B,
                                                                    images=images)
    img_b = images.download().reshape(B, 45, 128)
    for b in (0, 1, 2, 7, B - 1):
        row = api.Array.create(S)
        row.upload(h_b[b])
        img = api.Image.create(128, 45)
        api.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, s["gkd"], row, img)
        assert np.array_equal(img.download().view(np.uint32), img_b[b].view(np.uint32)), b
        i_ref = oracle.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, s["sv"], h_b[b])
        assert np.all(np.abs(img_b[b].astype(np.float64) - i_ref) <= 1e-5 * np.abs(h_b[b]).sum() + 1e-12)
    flat = img_b.reshape(B, -1)
    assert np.array_equal(vmin, flat.min(1)) and np.array_equal(vmax, flat.max(1))
    assert np.array_equal(imin, flat.argmin(1).astype(np.uint32)) and np.array_equal(imax, flat.argmax(1).astype(np.uint32))
    # without images, same extrema
    v2 = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, s["gkd"], hists, B)
    assert np.array_equal(v2[0], vmin) and np.array_equal(v2[3], imax)


def test_batched_k5_optimizer_inner_loop(api, oracle, setup):
    """Config 3: remove views -> K2 -> K3 -> candidate positions through batched K5 + K6 + arg-min
    (apps/optimize_trajectory/optimize_trajectory.cu:313-428)."""
    s = setup
    sc, S = s["sc"], s["sv"].shape[0]
    views = sc.views28()
    removed = views[[5, 40, 90]]
    obs_o = oracle.ObsRays(sc.num_points, 32)
    oracle.update_observation_rays(s["obvh"], True, views, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    oracle.update_observation_rays(s["obvh"], False, removed, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    obs = api.VectorArray.create(sc.num_points, 32)
    recons = api.Array.create(sc.num_points)
    api.update_observation_rays_batch(True, views, sc.max_distance, sc.width, sc.height, s["gbvh"], s["cloud"], obs)
    api.update_observation_rays_batch(False, removed, sc.max_distance, sc.width, sc.height, s["gbvh"], s["cloud"], obs)
    api.process_observation_rays(obs)
    api.evaluate_observation_rays(obs, recons)
    assert np.array_equal(obs.download(data=False)[1], obs_o.num_rows)
    # 7 Nelder-Mead candidates around each removed view
    rng = np.random.default_rng(4)
    cand = (removed[:, None, :3] + rng.normal(0, 2.0, (3, 7, 3))).reshape(-1, 3).astype(np.float32)
    B = cand.shape[0]
    hists = api.Array.create(B * S)
    hists.null()
    api.populate_spherical_histogram_batch(cand, sc.max_distance, s["gbvh"], s["cloud"], s["gkd"], hists,
                                           target_recon=sc.target_recon, obs_rays=obs, recons=recons)
    h_b = hists.download().reshape(B, S)
    r_g = recons.download()
    for b in range(B):
        h_ref, bins, deltas = oracle.populate_spherical_histogram_recon(
            s["obvh"], s["okd"], cand[b], sc.max_distance, sc.target_recon, sc.points, sc.normals, obs_o, r_g, S)
        assert _k5_close(h_b[b], h_ref, bins, deltas, S), b
    vmin, imin, vmax, imax = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, s["gkd"], hists, B)
    for b in (0, 9, B - 1):
        i_ref = oracle.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, s["sv"], h_b[b])
        assert abs(float(vmin[b]) - float(i_ref.min())) <= 1e-5 * np.abs(h_b[b]).sum() + 1e-12
    assert (vmin < 0).any()


@pytest.mark.parametrize("max_distance", [-7.5, 0.0])
def test_batched_k4_non_positive_max_distance(api, oracle, setup, max_distance):
    """`scale = 1 - l / max_distance` (kernels.cu:310-311) with max_distance <= 0: a negative value puts nothing out of
    range (the planner's swapped arguments pass a negative average there, plan_trajectory.cu:270-274), zero
    rejects every point at positive distance.  The conservative block culls must not interfere."""
    s = setup
    sc, S = s["sc"], s["sv"].shape[0]
    pos = np.array(POSITIONS[:3], np.float32)
    hists = api.Array.create(len(pos) * S)
    hists.null()
    api.populate_spherical_histogram_batch(pos, max_distance, s["gbvh"], s["cloud"], s["gkd"], hists)
    h_b = hists.download().reshape(len(pos), S)
    for b in range(len(pos)):
        h_ref, bins, scores = oracle.populate_spherical_histogram(s["obvh"], s["okd"], pos[b], max_distance,
                                                                  sc.points, sc.normals, s["values"], s["quals"], S)
        assert _hist_close(h_b[b], h_ref, bins, scores, S), b
        single = api.Array.create(S)
        single.null()
        api.populate_spherical_histogram(pos[b], max_distance, s["gbvh"], s["cloud"], s["gkd"], single)
        assert _hist_close(single.download(), h_ref, bins, scores, S), b
    assert (h_b != 0).any() == (max_distance < 0)


def test_certified_nn_lookup_equals_kd_search(api, oracle, setup):
    """The batched histogram path answers `find_nn` (libs/cacc/nnsearch.h:67-155) through a cube-map proposal certified
    against neighbour lists, with a precomputed rule for which of two coincident icosphere vertices the depth-first
    search meets first (kd.cuh).  It must return the SAME vertex index as the search, duplicates included."""
    s = setup
    kd, sv = s["gkd"], s["sv"].astype(np.float32)
    rng = np.random.default_rng(11)
    unit = lambda a: (a / np.linalg.norm(a, axis=1, keepdims=True)).astype(np.float32)
    rand = unit(rng.normal(size=(3_000_000, 3)))
    distinct = np.unique(sv, axis=0)
    d = np.linalg.norm(distinct[:, None, :] - distinct[None, :, :], axis=2) + np.eye(len(distinct)) * 9.0
    pairs = np.argwhere(d < 1.3 * d.min(axis=1, keepdims=True))                # neighbouring positions
    mid = 0.5 * (distinct[pairs[:, 0]] + distinct[pairs[:, 1]])
    bisector = unit(np.repeat(mid, 40, axis=0) + rng.normal(0, 2e-3, (len(mid) * 40, 3)))      # around the Voronoi edges
    bisector_tight = unit(np.repeat(mid, 20, axis=0) + rng.normal(0, 1e-6, (len(mid) * 20, 3)))
    at_vertex = np.concatenate([sv, unit(np.repeat(sv, 60, axis=0) + rng.normal(0, 5e-3, (len(sv) * 60, 3))),
                                unit(np.repeat(sv, 20, axis=0) + rng.normal(0, 1e-7, (len(sv) * 20, 3)))])   # tie-break planes
    axes = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1], [0, 0, 0],
                     [1, 1, 1], [-1, 1, -1], [1e-30, 0, 0], [1e30, 0, 0], [np.nan, 0, 1]], np.float32)
    scaled = np.concatenate([0.5 * rand[:200_000], 2.0 * rand[:200_000], 0.99 * rand[:200_000]])
    q = np.concatenate([rand, bisector, bisector_tight, at_vertex, axes, scaled]).astype(np.float32)
    plain = kd.find_nn(q, certified_lookup=False)
    fast = kd.find_nn(q, certified_lookup=True)
    assert np.array_equal(plain, fast)
    # both duplicates of coincident positions occur as answers (the rule is query dependent), and the device search is
    # the oracle's (reference acc::KDTree::find_nn restated) on a sample
    _, inv, cnt = np.unique(sv, axis=0, return_inverse=True, return_counts=True)
    inv = inv.reshape(-1)
    assert (cnt == 2).sum() == 630 and (cnt == 1).sum() == 12
    answered = np.zeros(len(sv), bool)
    answered[np.unique(fast[fast < len(sv)])] = True                   # the NaN query answers NAI, as in the search
    both = np.bincount(inv, weights=answered.astype(np.float64)) == 2   # positions whose two copies both win somewhere
    assert both.sum() >= 20 and np.bincount(inv, weights=answered.astype(np.float64)).min() >= 1
    sample = rng.choice(len(q) - len(axes) - len(scaled), 20_000, replace=False)
    assert np.array_equal(s["okd"].find_nn(q[sample])[0], fast[sample])


def test_certified_nn_lookup_refuses_unsuitable_sets(api, oracle):
    """Three coincident vertices have no precomputed tie rule: the lookup is not built, the plain search answers."""
    rng = np.random.default_rng(5)
    pts = rng.normal(size=(50, 3)).astype(np.float32)
    pts[7] = pts[3]; pts[9] = pts[3]
    okd = oracle.KD.build(pts)
    kn, kv = okd.export()
    kd = api.KDTree.create(kn, kv)
    q = rng.normal(size=(1000, 3)).astype(np.float32)
    assert np.array_equal(kd.find_nn(q), okd.find_nn(q)[0])
    with pytest.raises(RuntimeError):
        kd.find_nn(q, certified_lookup=True)
    # a generic set without triples: lookup available, same answers
    pts2 = rng.normal(size=(300, 3)).astype(np.float32)
    pts2[10] = pts2[20]
    okd2 = oracle.KD.build(pts2)
    kd2 = api.KDTree.create(*okd2.export())
    q2 = np.concatenate([rng.normal(size=(200_000, 3)), np.repeat(pts2, 50, axis=0) + rng.normal(0, 1e-3, (15000, 3))]).astype(np.float32)
    assert np.array_equal(kd2.find_nn(q2, certified_lookup=True), kd2.find_nn(q2))
    assert np.array_equal(kd2.find_nn(q2[:5000], certified_lookup=True), okd2.find_nn(q2[:5000])[0])
