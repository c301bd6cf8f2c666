"""The optimiser's simplex-downhill step on the device (uavmvs_nm_step) against the REFERENCE'S OWN nelder_mead()
(libs/opti/nelder_mead.h compiled into oracle/_ref/libref_geom.so) driven by the oracle's objective: altitude test,
acc closest_point airspace test, K5 + K6 + first arg-min of w * image (optimize_trajectory.cu:366-438)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _objective(oracle, ctx, track):
    sc, md = ctx["sc"], np.float32(ctx["min_distance"])

    def f(pos):
        pos = np.asarray(pos, np.float32)
        if pos[2] < md:
            ctx["early"] += 1
            return float(md - pos[2])
        w = np.float32(1.0)
        found, tri, cp = ctx["obvh"].closest_point(pos[None, :], float(md))
        if found[0]:
            t = ctx["tris"][tri[0]].reshape(3, 3)
            normal = np.cross((t[1] - t[0]).astype(np.float32), (t[2] - t[0]).astype(np.float32)).astype(np.float32)
            cpp = (pos - cp[0]).astype(np.float32)
            dist = np.float32(np.sqrt(np.float32(cpp[0] * cpp[0]) + np.float32(cpp[1] * cpp[1]) + np.float32(cpp[2] * cpp[2])))
            if float(np.dot(normal.astype(np.float64), cpp.astype(np.float64))) < 0.0:
                ctx["early"] += 1
                return float(dist)
            ctx["weighted"] += 1
            w = np.float32(1.0) - (md - dist) / md
        h, _, _ = oracle.populate_spherical_histogram_recon(ctx["obvh"], ctx["okd"], pos, sc.max_distance, sc.target_recon,
                                                            sc.points, sc.normals, ctx["obs_o"], ctx["recons"], ctx["S"])
        img = oracle.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, ctx["sv"], h.astype(np.float32)).reshape(-1)
        prod = (np.float32(w) * img.astype(np.float32)).astype(np.float32)
        mn, th, ph = 0.0, 0.0, 0.0
        if prod.min() < 0.0:
            k = int(prod.argmin())
            mn = float(prod[k])
            y, x = divmod(k, 128)
            th = float((np.float32(0.5) + (np.float32(y) / np.float32(45)) / np.float32(2)) * np.float32(3.14159265))
            ph = float((np.float32(x) / np.float32(128)) * np.float32(2) * np.float32(3.14159265))
        if mn < track["vmin"]:
            track.update(vmin=mn, vtheta=th, vphi=ph)
        return mn
    return f


def test_nelder_mead_step_equals_the_reference(oracle):
    if oracle.refgeom() is None or not hasattr(oracle.refgeom(), "ref_nelder_mead_step"):
        pytest.skip("oracle/_ref/libref_geom.so (reference nelder_mead.h) was not built")
    from uavmvs_b200 import api, scenes
    sc = scenes.dense()
    obvh = oracle.BVH.build(sc.verts, sc.faces, threads=2)
    arrays = obvh.export()
    gbvh = api.BVHTree.create(arrays.nodes, arrays.aabbs, arrays.tris, arrays.indices)
    sv = api.sphere_vertices(1.0, 3)
    okd = oracle.KD.build(sv)
    kn, kv = okd.export()
    gkd = api.KDTree.create(kn, kv)
    S = sv.shape[0]
    cloud = api.PointCloud.create(sc.points, sc.normals)
    views = sc.views28()
    obs = api.VectorArray.create(sc.num_points, 32)
    recons, wrecons = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
    api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height, gbvh, cloud, obs, recons, wrecons)
    obs_o = oracle.ObsRays(sc.num_points, 32)
    oracle.update_observation_rays(obvh, True, views, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    ctx = dict(sc=sc, obvh=obvh, okd=okd, sv=sv, S=S, obs_o=obs_o, recons=recons.download(), tris=arrays.tris, min_distance=4.0, early=0, weighted=0)
    # simplices: a tetrahedron around every 20th view, one lowered below the altitude limit, one pushed into a building
    sel = list(range(0, sc.num_views, 20))[:8]
    V = len(sel)
    rng = np.random.default_rng(5)
    base = views[sel, :3].copy()
    base[1, 2] = 5.0                                     # next to the ground: some vertices fail the altitude test
    base[2] = sc.points[100] + np.array([0.0, 0.0, 2.0], np.float32)     # within min_distance of the surface
    offs = np.array([[1, 1, 1], [1, -1, -1], [-1, 1, -1], [-1, -1, 1]], np.float32) * 1.5
    simplices = (base[:, None, :] + offs[None, :, :] + rng.normal(0, 0.2, (V, 4, 3))).astype(np.float32)
    nm = api.NelderMead.create(V, S)
    nm.upload(simplices)
    cur = simplices.copy()
    for step in range(3):
        nm.step(np.arange(V, dtype=np.uint32), ctx["min_distance"], sc.max_distance, sc.target_recon, sc.calib[0], sc.width, sc.height,
                gbvh, cloud, gkd, obs, recons)
        got = nm.download()
        for v in range(V):
            track = dict(vmin=0.0, vtheta=0.0, vphi=0.0)
            new, vid, val = oracle.ref_nelder_mead_step(cur[v], _objective(oracle, ctx, track))
            scale = np.abs(new).max()
            assert np.all(np.abs(got["simplices"][v] - new) <= 1e-5 * scale), (step, v)
            assert int(got["vid"][v]) == vid, (step, v)
            assert abs(float(got["value"][v]) - val) <= 1e-5 * max(abs(val), abs(float(got["value"][v]))) + 1e-6, (step, v)
            assert abs(float(got["vmin"][v]) - track["vmin"]) <= 1e-5 * abs(track["vmin"]) + 1e-6, (step, v)
            if track["vmin"] < -1e-3:                    # the orientation of a clear minimum
                assert abs(float(got["vtheta"][v]) - track["vtheta"]) < 1e-6 and abs(float(got["vphi"][v]) - track["vphi"]) < 1e-6, (step, v)
            cur[v] = new
    # the constraint branches of the objective were exercised (early returns, airspace weights) beside plain histogram values
    assert ctx["early"] > 0 and ctx["weighted"] > 0 and (got["value"] < 0).any()
