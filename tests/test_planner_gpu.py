"""Greedy planner (apps/plan_trajectory/plan_trajectory.cu:224-352) as a caller of the batched device path:
every step's 27 candidate scores against the oracle evaluating the candidates one at a time the way the
reference app does (K5, K6, host arg-max), then the adopted view through K1 + K3 on both sides."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from uavmvs_b200 import api as a
    assert a.device_count() >= 1
    return a


def _oracle_scores(oracle, obvh, okd, sv, sc, planner, pos, pen, ok, obs_o, recons_o, k5_args):
    scores = np.zeros(27, np.float32)
    scale = np.zeros(27)
    for j in np.nonzero(ok)[0]:
        h, bins, _ = oracle.populate_spherical_histogram_recon(obvh, okd, pos[j], k5_args[0], k5_args[1], sc.points,
                                                               sc.normals, obs_o, recons_o, sv.shape[0])
        img = oracle.evaluate_spherical_histogram(planner.calib, sc.width, sc.height, sv, h.astype(np.float32))
        mx = max(float(-img.min()), 0.0)                     # objective="improvement"
        scores[j] = np.float32(np.float32(mx) * planner.oweights[j]) * np.float32(1.0 - pen[j])
        scale[j] = 16.0 * max(int((bins != 0xFFFFFFFF).sum()), 1)
    return scores, scale


@pytest.mark.parametrize("reference_order", [False, True])
def test_planner_steps_match_oracle(api, oracle, reference_order):
    from uavmvs_b200 import planner as pl, scenes
    sc = scenes.dense()
    obvh = oracle.BVH.build(sc.verts, sc.faces, threads=2)
    ta = obvh.export()
    gbvh = api.BVHTree.create(ta.nodes, ta.aabbs, ta.tris, ta.indices)
    sv = oracle.sphere_vertices(1.0, 3)
    okd = oracle.KD.build(sv)
    kn, kv = okd.export()
    gkd = api.KDTree.create(kn, kv)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    p = pl.Planner(gbvh, cloud, gkd, sc.num_points, sv.shape[0], cloud_points=sc.points, min_distance=1.0,
                   max_distance=sc.max_distance, max_altitude=200.0, max_velocity=8.0, width=sc.width, height=sc.height,
                   start_pos=(20.0, 24.0, 30.0), start_vel=(3.0, 0.5, 0.0), reference_argument_order=reference_order,
                   objective="improvement")
    # a fresh state has no rows: K5 adds nothing (`if (num_rows >= 1)`, kernels.cu:378), so seed two views first
    seed = sc.views28()[[0, sc.num_views // 2]]
    api.update_observation_rays_batch(True, seed, sc.max_distance, sc.width, sc.height, gbvh, cloud, p.obs)
    api.evaluate_observation_rays(p.obs, p.recons)
    if reference_order:
        p.avg_recon = np.float32(45.0)       # the swapped argument is the K5 range: start with something in reach
    obs_o = oracle.ObsRays(sc.num_points, 20)
    oracle.update_observation_rays(obvh, True, seed, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    recons_o = oracle.evaluate_observation_rays(obs_o)
    steps = 0
    for it in range(5):
        pos, pen, ok = p.candidates()
        assert ok.sum() >= 9
        k5 = (p.avg_recon, p.max_distance) if reference_order else (p.max_distance, p.avg_recon)
        s_o, scale = _oracle_scores(oracle, obvh, okd, sv, sc, p, pos, pen, ok, obs_o, recons_o, k5)
        s_g, rots = p.score(pos, pen, ok)
        assert np.all(np.abs(s_g.astype(np.float64) - s_o) <= 1e-5 * scale + 1e-6), (it, s_g, s_o)
        j = int(np.argmax(s_g))
        top = np.sort(s_o)[::-1]
        if top[0] - top[1] > 2e-5 * scale.max():                 # clear winner: same decision
            assert j == int(np.argmax(s_o))
        if s_g[j] < 0.1:
            break
        avg_g = p.advance(j, pos, rots)
        # the adopted view on the oracle side: K1 populate + K3 (plan_trajectory.cu:326-337)
        from uavmvs_b200 import formats
        view = formats.views_from_trajectory(pos[j][None], rots[j][None], [p.flen], sc.width, sc.height)
        oracle.update_observation_rays(obvh, True, view, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
        recons_o = oracle.evaluate_observation_rays(obs_o)
        assert np.array_equal(p.obs.download(data=False)[1], obs_o.num_rows)
        assert np.allclose(p.recons.download(), recons_o, rtol=1e-5, atol=1e-7)
        assert abs(float(avg_g) - float(recons_o.astype(np.float64).mean())) <= 1e-5
        steps += 1
    assert steps >= 3
    assert len(p.trajectory) == steps
    rot = p.trajectory[-1][1]
    assert np.allclose(rot @ rot.T, np.eye(3), atol=1e-5)        # rotation_from_spherical is orthonormal


def test_plan_trajectory_writes_a_reference_trajectory_file(api, oracle, tmp_path):
    from uavmvs_b200 import planner as pl, scenes, formats
    sc = scenes.tiny()
    gbvh = api.BVHTree.build_sah(sc.verts, sc.faces)
    sv = oracle.sphere_vertices(1.0, 3)
    kn, kv = oracle.KD.build(sv).export()
    gkd = api.KDTree.create(kn, kv)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    c = sc.points.mean(0)
    pos, rot, flen = pl.plan_trajectory(gbvh, cloud, gkd, sc.num_points, sv.shape[0], num_views=6,
                                        max_distance=sc.max_distance, max_altitude=500.0, max_velocity=10.0,
                                        width=sc.width, height=sc.height, start_pos=(c[0], c[1], c[2] + 25.0),
                                        start_vel=(2.0, 0.0, 0.0), reference_argument_order=False)
    assert len(pos) == 0       # the reference's objective never fires (planner.py docstring): empty plan, like the app
    p = pl.Planner(gbvh, cloud, gkd, sc.num_points, sv.shape[0], max_distance=sc.max_distance, max_altitude=500.0,
                   max_velocity=10.0, width=sc.width, height=sc.height, start_pos=(c[0], c[1], c[2] + 25.0),
                   start_vel=(2.0, 0.0, 0.0), reference_argument_order=False, objective="improvement")
    api.update_observation_rays_batch(True, sc.views28()[:2], sc.max_distance, sc.width, sc.height, gbvh, cloud, p.obs)
    api.evaluate_observation_rays(p.obs, p.recons)
    for _ in range(6):
        if not p.step():
            break
    assert 1 <= len(p.trajectory) <= 6
    pos = np.array([t[0] for t in p.trajectory], np.float32)
    rot = np.array([t[1] for t in p.trajectory], np.float32)
    flen = np.full(len(pos), p.flen, np.float32)
    path = str(tmp_path / "plan.traj")
    formats.save_trajectory(path, pos, rot, flen)
    p2, r2, f2 = formats.load_trajectory(path)
    assert np.allclose(p2, pos, atol=1e-4) and np.allclose(r2, rot, atol=1e-5)
