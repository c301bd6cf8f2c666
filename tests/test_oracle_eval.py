"""The restated kernel math (libs/eval/kernels.cu) of the oracle: frozen fixture, closed-form
known answers and the properties the reference's apps rely on."""
import math
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PI = np.float32(3.14159265)


def _eval(oracle, sc, obvh, views=None, max_rows=32):
    obs = oracle.ObsRays(sc.num_points, max_rows)
    v = sc.views28() if views is None else views
    state, cnt = oracle.update_observation_rays(obvh, True, v, sc.max_distance, sc.width, sc.height,
                                                sc.points, sc.normals, obs, want_state=True, want_counters=True)
    return obs, state, cnt


def test_frozen_fixture(oracle, tiny_scene, tiny_tree):
    import hashlib
    sha = lambda a: np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)  # noqa: E731
    g = np.load(os.path.join(GOLD, "port_eval_tiny.npz"))
    sc = tiny_scene
    obvh, _ = tiny_tree
    obs, state, cnt = _eval(oracle, sc, obvh)
    assert np.array_equal(np.packbits(state == 2), g["state"]) and np.array_equal(np.packbits(state >= 1), g["traced"])
    assert np.array_equal(cnt, g["counters"])
    assert np.array_equal(obs.num_rows, g["rows_k1"]) and np.array_equal(sha(obs.data), g["data_k1_sha"])
    oracle.process_observation_rays(obs)
    assert np.array_equal(obs.num_rows, g["rows"]) and np.array_equal(sha(obs.data), g["data_sorted_sha"])
    recons = oracle.evaluate_observation_rays(obs)
    assert np.array_equal(recons.view(np.uint32), g["recons"].view(np.uint32))
    wrecons = oracle.calculate_func_recons(recons, sc.target_recon)
    assert np.array_equal(wrecons.view(np.uint32), g["wrecons"].view(np.uint32))
    sv = oracle.sphere_vertices(1.0, 3)
    kd = oracle.KD.build(sv)
    h4, _, _ = oracle.populate_spherical_histogram(obvh, kd, g["hist_pos"], sc.max_distance, sc.points, sc.normals,
                                                   g["values"], g["qualities"], sv.shape[0])
    h5, _, _ = oracle.populate_spherical_histogram_recon(obvh, kd, g["hist_pos"], sc.max_distance, sc.target_recon,
                                                         sc.points, sc.normals, obs, recons, sv.shape[0])
    assert np.array_equal(h4, g["hist4"]) and np.array_equal(h5, g["hist5"])
    img = oracle.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, sv, h5.astype(np.float32))
    assert np.array_equal(img.view(np.uint32), g["image5"].view(np.uint32))


def _one_triangle_scene():
    """A ground triangle, one point on it, cameras overhead (hand-checkable)."""
    from uavmvs_b200 import scenes
    verts = np.array([[-50, -50, 0], [50, -50, 0], [0, 50, 0], [-1, -1, 5], [1, -1, 5], [0, 1, 5]], np.float32)
    faces = np.array([[0, 1, 2], [3, 4, 5]], np.uint32)      # big ground + a small occluder 5 m above the origin
    return scenes, verts, faces


def test_known_answer_visibility(oracle):
    scenes, verts, faces = _one_triangle_scene()
    obvh = oracle.BVH.build(verts, faces, threads=1)
    pts = np.array([[0, 0, 0], [10, 0, 0], [0, 0, 0]], np.float32)
    nrm = np.array([[0, 0, 1], [0, 0, 1], [0, 0, -1]], np.float32)     # third point faces away
    specs = [((0.0, 0.0, 40.0), 0.0, 0.0),       # straight above the origin: occluder blocks point 0, sees point 1
             ((30.0, 0.0, 40.0), math.pi, math.atan2(30.0, 40.0)),   # oblique, aimed at the origin: sees both
             ((0.0, 0.0, 100.0), 0.0, 0.0),      # beyond max_distance = 80
             ((10.0, 0.0, 0.5), 0.0, 0.0)]       # grazing: cos(theta) < 0.087 for point 0, straight above point 1
    pos, w2c, calib = scenes._pack_views(specs)
    views = oracle.pack_views(pos, w2c, calib)
    obs = oracle.ObsRays(3, 32)
    state, cnt = oracle.update_observation_rays(obvh, True, views, 80.0, 1920, 1080, pts, nrm, obs,
                                                want_state=True, want_counters=True)
    assert state[:, 0].tolist() == [1, 2, 0, 0]      # 1 = traced and occluded, 2 = visible, 0 = culled
    assert state[:, 1].tolist() == [2, 2, 0, 2]
    assert state[:, 2].tolist() == [0, 0, 0, 0]
    assert obs.num_rows.tolist() == [1, 3, 0] and int(cnt[0]) == 5
    # stored ray of point 1 for view 3 (camera 0.5 m straight above it): direction +z, scale 1 - 0.5/80
    r = obs.data[2, 1]
    assert abs(r[2] - 1.0) < 1e-6 and abs(r[0]) < 1e-6 and abs(r[1]) < 1e-6
    assert r[3] == np.float32(1.0) - np.float32(0.5) / np.float32(80.0)


def test_known_answer_heuristic_pair(oracle):
    """Two observations at known parallax: recon = m(alpha) * t(alpha) in closed form
    (libs/eval/kernels.cu:111-136)."""
    for alpha_deg in (5.0, 20.0, 45.0, 90.0):
        a = math.radians(alpha_deg)
        r0 = np.array([math.sin(a / 2), 0.0, math.cos(a / 2), 0.8], np.float32)
        r1 = np.array([-math.sin(a / 2), 0.0, math.cos(a / 2), 0.6], np.float32)
        obs = oracle.ObsRays(1, 32)
        obs.data[0, 0], obs.data[1, 0] = r0, r1
        obs.num_rows[0] = 2
        recon = float(oracle.evaluate_observation_rays(obs)[0])
        L = lambda x, k, x0: 1.0 / (1.0 + math.exp(-k * (x - x0)))  # noqa: E731
        alpha = math.acos(float(np.float32(r0[0] * r1[0]) + np.float32(r0[2] * r1[2])))
        expect = (1.0 - L(alpha, 8.0, math.pi / 4.0)) * min(r0[2], r1[2]) * L(alpha, 32.0, math.pi / 16.0) * 0.6
        assert abs(recon - float(expect)) <= 1e-5 * abs(float(expect)), (alpha_deg, recon, expect)
    obs = oracle.ObsRays(2, 32)
    obs.data[0, 1] = (0, 0, 1, 0.5)
    obs.num_rows[:] = (0, 1)
    assert oracle.evaluate_observation_rays(obs).tolist() == [-1.0, 0.0]      # kernels.cu:283-288
    w = oracle.calculate_func_recons(np.array([-1.0, 0.0, 2.5, 3.0, 7.0], np.float32), 3.0)
    assert np.allclose(w, [16.0, 9.0, 0.25, 0.0, 0.0])


def test_sort_is_descending_and_stable_under_removal(oracle, tiny_tree, tiny_scene):
    """K2 orders by cos(theta) descending, invalidated rows leave (kernels.cu:204-264)."""
    from uavmvs_b200 import scenes
    sc = scenes.dense()
    obvh = oracle.BVH.build(sc.verts, sc.faces, threads=2)
    obs, _, _ = _eval(oracle, sc, obvh)
    rows_before = obs.num_rows.copy()
    oracle.process_observation_rays(obs)
    assert np.array_equal(obs.num_rows, rows_before)           # nothing invalid: counters untouched (:251-254)
    r = np.minimum(obs.num_rows, 32)
    z = obs.data[:, :, 2]
    for i in np.flatnonzero(r >= 2)[:500]:
        col = z[:r[i], i]
        assert np.all(col[:-1] >= col[1:])
    removed = sc.views28()[[5, 50, 90]]
    oracle.update_observation_rays(obvh, False, removed, sc.max_distance, sc.width, sc.height, sc.points,
                                   sc.normals, obs)
    oracle.process_observation_rays(obs)
    # removing views and re-evaluating equals evaluating without them (where no column overflowed)
    keep = np.ones(sc.num_views, bool)
    keep[[5, 50, 90]] = False
    obs2, _, _ = _eval(oracle, sc, obvh, views=sc.views28()[keep])
    oracle.process_observation_rays(obs2)
    ok = rows_before <= 32
    assert np.array_equal(obs.num_rows[ok], obs2.num_rows[ok])
    a, b = oracle.evaluate_observation_rays(obs), oracle.evaluate_observation_rays(obs2)
    assert np.allclose(a[ok], b[ok], rtol=1e-5, atol=1e-6)


def test_histogram_image_counts_bins_in_frustum(oracle):
    """K6 with a uniform histogram counts the sphere vertices inside each orientation's frustum;
    it is invariant to which twin of a coincident pair received the mass."""
    from uavmvs_b200 import scenes
    sv = oracle.sphere_vertices(1.0, 3)
    _, K = scenes.make_view(np.zeros(3), 0.0, 0.0)
    ones = np.ones(sv.shape[0], np.float32)
    img = oracle.evaluate_spherical_histogram(K.astype(np.float32), 1920, 1080, sv, ones)
    assert img.shape == (45, 128) and img.min() >= 0 and img.max() > 10
    assert np.all(img == np.round(img))
    # move all mass of every duplicate onto the first twin: image unchanged
    _, first, inv = np.unique(sv, axis=0, return_index=True, return_inverse=True)
    merged = np.zeros_like(ones)
    np.add.at(merged, first[inv.reshape(-1)], ones)
    img2 = oracle.evaluate_spherical_histogram(K.astype(np.float32), 1920, 1080, sv, merged)
    assert np.array_equal(img, img2)


def test_oracle_sampler_known_answers(oracle):
    """generate_proxy_cloud.cpp:84-184 restated: one unit-area triangle at density d yields floor(d) or floor(d) + 1
    samples, all inside the triangle, with its normal; two identical meshes give identical clouds (seed = 3 * face)."""
    verts = np.array([[0, 0, 0], [2, 0, 0], [0, 1, 0], [10, 0, 0], [12, 0, 0], [10, 1, 0]], np.float32)
    faces = np.array([[0, 1, 2], [3, 4, 5]], np.uint32)
    p, n = oracle.sample_proxy_cloud(verts, faces, 40.5)
    assert 80 <= len(p) <= 82
    assert np.array_equal(n, np.tile(np.array([0, 0, 1], np.float32), (len(p), 1)))
    left = p[p[:, 0] < 5]
    assert np.all(left[:, 0] >= 0) and np.all(left[:, 1] >= 0) and np.all(left[:, 0] / 2 + left[:, 1] <= 1 + 1e-6)
    assert np.all(p[:, 2] == 0)
    p2, _ = oracle.sample_proxy_cloud(verts, faces, 40.5)
    assert np.array_equal(p, p2)
    # the Z-order key (libs/acc/math.h:19-45) is monotone in x on a line: samples with y = z = 0 come out x-sorted
    line = np.array([[0, 0, 0], [1, 0, 0], [0, 1e-9, 0]], np.float32)
    q, _ = oracle.sample_proxy_cloud(line, np.array([[0, 1, 2]], np.uint32), 4e10)
    assert len(q) > 5
