"""Parity of the CUDA path (through the C ABI) against the CPU oracle.

Bars (BASELINE.json north_star): visibility flags and observation counts bit-exact;
stored observation rays bit-exact (same operation order); heuristic scores within
1e-5 relative (libm vs CUDA acosf/expf).
"""
import numpy as np
import pytest

from conftest import make_rays

pytestmark = pytest.mark.gpu

REL = 1e-5


@pytest.fixture(scope="module")
def api():
    from uavmvs_b200 import api as a
    assert a.device_count() >= 1
    return a


def _close(a, b, rel=REL, floor=0.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.all(np.abs(a - b) <= rel * np.maximum(np.abs(a), np.abs(b)) + floor)


def _gpu_tree(api, arrays):
    return api.BVHTree.create(arrays.nodes, arrays.aabbs, arrays.tris, arrays.indices)


def test_trace_matches_oracle(api, oracle, tiny_scene, tiny_tree):
    obvh, arrays = tiny_tree
    gbvh = _gpu_tree(api, arrays)
    rays = make_rays(tiny_scene, 100_000, seed=1)
    h_o, f_o, t_o = obvh.intersect(rays, any_hit=False)
    h_g, f_g, t_g = gbvh.trace(rays, any_hit=False)
    assert np.array_equal(h_o, h_g)
    assert np.array_equal(f_o, f_g)
    assert np.array_equal(t_o.view(np.uint32), t_g.view(np.uint32))
    h_a, _, _ = gbvh.trace(rays, any_hit=True)
    assert np.array_equal(h_o, h_a)
    assert 0.05 < h_o.mean() < 0.95


def test_trace_degenerate_inputs(api, oracle):
    """Single-triangle tree (root is a leaf), a degenerate triangle, axis-parallel rays."""
    verts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [2, 2, 2], [2, 2, 2], [2, 2, 2]], np.float32)
    for faces in (np.array([[0, 1, 2]], np.uint32), np.array([[0, 1, 2], [3, 4, 5]], np.uint32)):
        obvh = oracle.BVH.build(verts, faces, threads=1)
        gbvh = _gpu_tree(api, obvh.export())
        rays = np.array([[0.2, 0.2, 1, 0, 0, -1, 0.001, 5.0],
                         [0.2, 0.2, 1, 0, 0, 1, 0.001, 5.0],
                         [1.0005, 0.0, 1, 0, 0, -1, 0.0, 5.0],     # inside the 1e-3 slack
                         [1.01, 0.0, 1, 0, 0, -1, 0.0, 5.0],       # outside
                         [2, 2, 3, 0, 0, -1, 0.0, 5.0]], np.float32)
        for any_hit in (True, False):
            h_o, f_o, _ = obvh.intersect(rays, any_hit=any_hit)
            h_g, f_g, _ = gbvh.trace(rays, any_hit=any_hit)
            assert np.array_equal(h_o, h_g)
            if not any_hit:
                assert np.array_equal(f_o, f_g)
            # one triangle: root is a leaf, no box test, the slack-zone ray hits; two faces: the
            # leaf's tight box rejects it first (tree-dependent visibility)
            assert list(h_o[:4]) == ([1, 0, 1, 0] if len(faces) == 1 else [1, 0, 0, 0])


def _run_oracle_eval(oracle, obvh, sc, views=None, max_rows=32):
    obs = oracle.ObsRays(sc.num_points, max_rows)
    v = sc.views28() if views is None else views
    state, cnt = oracle.update_observation_rays(obvh, True, v, sc.max_distance, sc.width, sc.height,
                                                sc.points, sc.normals, obs, want_state=True, want_counters=True)
    return obs, state, cnt


def _valid_rows_equal(d_o, d_g, rows, max_rows):
    r = np.minimum(rows, max_rows)
    mask = np.arange(max_rows)[:, None] < r[None, :]
    return np.array_equal(d_o.view(np.uint32)[mask], d_g.view(np.uint32)[mask])


def test_update_observation_rays_batch(api, oracle, tiny_scene, tiny_tree):
    sc = tiny_scene
    obvh, arrays = tiny_tree
    obs_o, state, cnt = _run_oracle_eval(oracle, obvh, sc)
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, sc.views28(), sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    data, rows = obs.download()
    assert np.array_equal(rows, obs_o.num_rows)                 # observation counts: bit-exact
    assert rows.sum() == (state == 2).sum() and rows.sum() > 1000
    assert _valid_rows_equal(obs_o.data, data, rows, 32)          # stored rays: bit-exact


def test_single_view_calls_equal_batch(api, oracle, tiny_scene, tiny_tree):
    """The per-view entry point (reference launch loop) gives the same rows as the batch."""
    sc = tiny_scene
    obvh, arrays = tiny_tree
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    for k in range(sc.num_views):
        api.update_observation_rays(True, sc.view_pos[k], sc.max_distance, sc.w2c[k], sc.calib[k],
                                    sc.width, sc.height, gbvh, cloud, obs)
    data, rows = obs.download()
    obs_o, _, _ = _run_oracle_eval(oracle, obvh, sc)
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)


def test_row_cap_and_overflow_counter(api, oracle, tiny_scene, tiny_tree):
    """max_rows = 4: counters keep counting, rows beyond the cap are dropped (kernels.cu:176-178)."""
    sc = tiny_scene
    obvh, arrays = tiny_tree
    obs_o, _, _ = _run_oracle_eval(oracle, obvh, sc, max_rows=4)
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 4)
    api.update_observation_rays_batch(True, sc.views28(), sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    data, rows = obs.download()
    assert rows.max() > 4
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 4)


def test_process_and_evaluate(api, oracle, tiny_scene, tiny_tree):
    sc = tiny_scene
    obvh, arrays = tiny_tree
    obs_o, _, _ = _run_oracle_eval(oracle, obvh, sc)
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, sc.views28(), sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    api.process_observation_rays(obs)
    oracle.process_observation_rays(obs_o)
    data, rows = obs.download()
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)          # same sort network -> same order
    recons, wrecons = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
    api.evaluate_observation_rays(obs, recons)
    api.calculate_func_recons(recons, sc.target_recon, wrecons)
    r_o = oracle.evaluate_observation_rays(obs_o)
    w_o = oracle.calculate_func_recons(r_o, sc.target_recon)
    r_g, w_g = recons.download(), wrecons.download()
    assert np.array_equal(r_o == -1.0, r_g == -1.0)
    assert _close(r_o, r_g)
    assert _close(w_o, w_g, floor=1e-6)
    s, mn, mx = api.reduce(wrecons)
    ref = oracle.reduce(w_o)
    # sum / min / max of scores: 1e-5 relative like the scores themselves (wrecon is a difference: floor)
    assert _close(s, ref[0]) and _close(mn, ref[1], floor=1e-6) and _close(mx, ref[2], floor=1e-6)


def test_remove_views(api, oracle, tiny_scene, tiny_tree):
    """populate=false marks rays, K2 compacts (optimize_trajectory.cu:313-354)."""
    sc = tiny_scene
    obvh, arrays = tiny_tree
    views = sc.views28()
    removed = views[[3, 11, 17]]
    obs_o, _, _ = _run_oracle_eval(oracle, obvh, sc)
    oracle.update_observation_rays(obvh, False, removed, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, views, sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    before = obs.download(data=False)[1].sum()
    api.update_observation_rays_batch(False, removed, sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    api.process_observation_rays(obs)
    data, rows = obs.download()
    assert rows.sum() < before
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)


def test_reevaluate_only_changed_columns(api, oracle, tiny_scene, tiny_tree):
    """uavmvs_reevaluate_observation_rays = process + evaluate + calculate_func_recons on the blocks K1 touched
    (the optimiser's per-iteration re-evaluation, optimize_trajectory.cu:313-345, :499-529): identical to the
    three full launches after every step of remove -> re-evaluate -> re-add elsewhere -> re-evaluate."""
    sc = tiny_scene
    obvh, arrays = tiny_tree
    views = sc.views28()
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    n = sc.num_points
    obs, obs_full = api.VectorArray.create(n, 32), api.VectorArray.create(n, 32)
    rec, wrec = api.Array.create(n), api.Array.create(n)
    rec_f, wrec_f = api.Array.create(n), api.Array.create(n)
    obs_o = oracle.ObsRays(n, 32)

    def step(populate, v):
        oracle.update_observation_rays(obvh, populate, v, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
        oracle.process_observation_rays(obs_o)
        for o in (obs, obs_full):
            api.update_observation_rays_batch(populate, v, sc.max_distance, sc.width, sc.height, gbvh, cloud, o)
        api.reevaluate_observation_rays(obs, rec, sc.target_recon, wrec)
        api.process_observation_rays(obs_full)
        api.evaluate_observation_rays(obs_full, rec_f)
        api.calculate_func_recons(rec_f, sc.target_recon, wrec_f)
        d_i, r_i = obs.download()
        d_f, r_f = obs_full.download()
        assert np.array_equal(r_i, r_f) and np.array_equal(r_i, obs_o.num_rows)
        assert _valid_rows_equal(d_f, d_i, r_i, 32) and _valid_rows_equal(obs_o.data, d_i, r_i, 32)
        assert np.array_equal(rec.download().view(np.uint32), rec_f.download().view(np.uint32))     # same kernel, same bits
        assert np.array_equal(wrec.download().view(np.uint32), wrec_f.download().view(np.uint32))
        assert _close(rec.download(), oracle.evaluate_observation_rays(obs_o))

    step(True, views)                       # first call: every block is flagged
    step(False, views[[3, 11]])             # two views removed: few blocks flagged
    moved = views[[3, 11]].copy()
    moved[:, 0:3] += np.float32(1.5)
    for k in range(2):                      # w2c translation column for the moved centre
        w2c = moved[k, 3:19].reshape(4, 4).copy()
        w2c[:3, 3] = -(w2c[:3, :3].astype(np.float64) @ moved[k, :3].astype(np.float64))
        moved[k, 3:19] = w2c.reshape(16)
    step(True, moved)
    # nothing changed: a further call touches nothing and keeps the arrays
    before = rec.download().copy()
    api.reevaluate_observation_rays(obs, rec, sc.target_recon, wrec)
    assert np.array_equal(before.view(np.uint32), rec.download().view(np.uint32))
    # poisoned result arrays stay poisoned where no block is flagged, are rewritten after mark_dirty
    rec.upload(np.full(n, 123.0, np.float32))
    api.reevaluate_observation_rays(obs, rec, sc.target_recon, wrec)
    assert np.all(rec.download() == 123.0)
    obs.mark_dirty()
    api.reevaluate_observation_rays(obs, rec, sc.target_recon, wrec)
    assert np.array_equal(before.view(np.uint32), rec.download().view(np.uint32))


def test_evaluate_trajectory_fused(api, oracle, tiny_scene, tiny_tree):
    sc = tiny_scene
    obvh, arrays = tiny_tree
    obs_o, state, cnt = _run_oracle_eval(oracle, obvh, sc)
    oracle.process_observation_rays(obs_o)
    r_o = oracle.evaluate_observation_rays(obs_o)
    w_o = oracle.calculate_func_recons(r_o, sc.target_recon)
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    recons, wrecons = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
    for views in (sc.views28(), api.DeviceViews(sc.views28())):
        s, mn, mx, stats = api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height,
                                                   gbvh, cloud, obs, recons, wrecons)
        assert stats[0] == int(cnt[0])                         # occlusion rays traced
        assert stats[1] == int((state == 2).sum())             # visible observations
        assert np.array_equal(obs.download(data=False)[1], obs_o.num_rows)
        assert _close(recons.download(), r_o)
        assert _close(s / sc.num_points, w_o.astype(np.float64).mean())
    # the same evaluation with the per-point downloads overlapped (evaluate_trajectory.cu:188-194)
    r_dev, rows_dev = recons.download(), obs.download(data=False)[1]
    h_r, h_n = np.full(sc.num_points, -7.0, np.float32), np.full(sc.num_points, 77, np.uint32)
    s2, mn2, mx2, stats2 = api.evaluate_trajectory_to_host(sc.views28(), sc.max_distance, sc.target_recon, sc.width,
                                                           sc.height, gbvh, cloud, obs, recons, wrecons, h_r, h_n)
    assert stats2 == stats and (mn2, mx2) == (mn, mx) and _close(s2, s, 1e-6)
    assert np.array_equal(h_r.view(np.uint32), r_dev.view(np.uint32)) and np.array_equal(h_n, rows_dev)
    assert np.array_equal(recons.download().view(np.uint32), r_dev.view(np.uint32))


@pytest.mark.parametrize("builder", ["build_lbvh", "build_ploc", "build_sah"])
def test_gpu_built_tree_same_tree_parity(api, oracle, tiny_scene, builder):
    """GPU-built trees (LBVH, PLOC, sweep SAH): the oracle walks the exported tree (parity is per tree)."""
    sc = tiny_scene
    gbvh = getattr(api.BVHTree, builder)(sc.verts, sc.faces)
    nodes, aabbs, tris, indices = gbvh.export()
    n, t = gbvh.sizes()
    assert t == sc.num_tris
    if builder == "build_sah":
        # multi-triangle leaves like the reference's builder; every leaf range is covered exactly once
        leaf = nodes[:, 2] == 0xFFFFFFFF
        assert n % 2 == 1 and n < 2 * t - 1 and leaf.sum() == (n + 1) // 2
        order = np.argsort(nodes[leaf, 0])
        first, last = nodes[leaf, 0][order], nodes[leaf, 1][order]
        assert first[0] == 0 and last[-1] == t and np.array_equal(first[1:], last[:-1])
        assert (last - first).max() <= 8
    else:
        assert n == 2 * t - 1
    assert sorted(indices.tolist()) == list(range(t))
    # every inner box is the exact union of its children
    inner = nodes[:, 2] != 0xFFFFFFFF
    l, r = nodes[inner, 2], nodes[inner, 3]
    assert np.array_equal(aabbs[inner, :3], np.minimum(aabbs[l, :3], aabbs[r, :3]))
    assert np.array_equal(aabbs[inner, 3:], np.maximum(aabbs[l, 3:], aabbs[r, 3:]))
    # a proper binary tree: every node except the root has exactly one parent, the root is node 0
    kids = np.concatenate([l, r])
    assert len(np.unique(kids)) == len(kids) == n - 1 and 0 not in kids
    obvh = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    rays = make_rays(sc, 50_000, seed=2)
    h_o, f_o, t_o = obvh.intersect(rays, any_hit=False)
    h_g, f_g, t_g = gbvh.trace(rays, any_hit=False)
    assert np.array_equal(h_o, h_g) and np.array_equal(f_o, f_g)
    assert np.array_equal(gbvh.trace(rays, any_hit=True)[0], h_o)
    obs_o, state, _ = _run_oracle_eval(oracle, obvh, sc)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, sc.views28(), sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    data, rows = obs.download()
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)


def _populate_both(api, oracle, sc, arrays):
    obvh = oracle.BVH.from_arrays(arrays)
    obs_o, state, _ = _run_oracle_eval(oracle, obvh, sc)
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, sc.views28(), sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    data, rows = obs.download()
    return obs_o, data, rows, state


def test_dense_scene_rows_and_cap(api, oracle):
    """~15 observations per point, some beyond the 32-row cap: row order, cap, un-clamped counters."""
    from uavmvs_b200 import scenes
    sc = scenes.dense()
    arrays = oracle.BVH.build(sc.verts, sc.faces, threads=2).export()
    obs_o, data, rows, state = _populate_both(api, oracle, sc, arrays)
    assert rows.max() > 32 and rows.mean() > 8
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)


def test_non_nested_tree_takes_exact_path(api, oracle, tiny_scene, tiny_tree):
    """User-supplied boxes that are not exact unions (here: some boxes SHRUNK, so rays the tight tree
    would accept are rejected): the answer is tree dependent and must still equal the reference walk."""
    sc = tiny_scene
    _, arrays = tiny_tree
    rng = np.random.default_rng(4)
    aabbs = arrays.aabbs.copy()
    pick = rng.random(aabbs.shape[0]) < 0.2
    pick[0] = False
    ext = aabbs[:, 3:] - aabbs[:, :3]
    aabbs[pick, :3] += 0.2 * ext[pick]
    aabbs[pick, 3:] -= 0.2 * ext[pick]
    mod = oracle.TreeArrays(arrays.nodes, aabbs, arrays.tris, arrays.indices)
    obs_o, data, rows, state = _populate_both(api, oracle, sc, mod)
    obs_t, _, _ = _run_oracle_eval(oracle, oracle.BVH.from_arrays(arrays), sc)
    assert not np.array_equal(obs_t.num_rows, obs_o.num_rows)      # the shrunk boxes do change visibility
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)


def test_large_leaves(api, oracle, tiny_scene):
    """A hand-made tree whose leaves hold more than 31 triangles."""
    sc = tiny_scene
    t = sc.verts[sc.faces].reshape(-1, 9).astype(np.float32)
    order = np.argsort(t[:, 0::3].mean(1), kind="stable")       # split along x into 2^k slabs
    t, idx = t[order], order.astype(np.uint32)
    nodes, aabbs = [], []

    def box(lo, hi):
        v = t[lo:hi].reshape(-1, 3)
        return np.concatenate([v.min(0), v.max(0)])

    def build(lo, hi, depth):
        me = len(nodes)
        nodes.append([lo, hi, 0xFFFFFFFF, 0xFFFFFFFF])
        aabbs.append(box(lo, hi))
        if depth < 5 and hi - lo > 80:
            mid = (lo + hi) // 2
            l = build(lo, mid, depth + 1)
            r = build(mid, hi, depth + 1)
            nodes[me][2], nodes[me][3] = l, r
        return me

    build(0, t.shape[0], 0)
    arrays = oracle.TreeArrays(np.array(nodes, np.uint32), np.array(aabbs, np.float32), t, idx)
    assert (arrays.nodes[arrays.nodes[:, 2] == 0xFFFFFFFF, 1] - arrays.nodes[arrays.nodes[:, 2] == 0xFFFFFFFF, 0]).max() > 31
    obs_o, data, rows, state = _populate_both(api, oracle, sc, arrays)
    assert np.array_equal(rows, obs_o.num_rows) and rows.sum() > 1000
    assert _valid_rows_equal(obs_o.data, data, rows, 32)


def test_axis_parallel_rays_and_flat_boxes(api, oracle):
    """Cameras exactly above points (zero direction components) over axis-aligned flat geometry
    (zero-thickness leaf boxes): the cases the filtered slab test hands to the exact path."""
    from uavmvs_b200 import scenes
    g = np.arange(0, 9, dtype=np.float32)
    X, Y = np.meshgrid(g, g, indexing="ij")
    gv = np.stack([X.ravel(), Y.ravel(), np.zeros(81, np.float32)], 1)
    quads = [(i * 9 + j, (i + 1) * 9 + j, (i + 1) * 9 + j + 1, i * 9 + j + 1) for i in range(8) for j in range(8)]
    gf = np.array([[a, b, c] for a, b, c, d in quads] + [[a, c, d] for a, b, c, d in quads], np.uint32)
    # a flat roof plate 3 m above part of the ground
    rv = np.array([[2, 2, 3], [5, 2, 3], [5, 5, 3], [2, 5, 3]], np.float32)
    rf = np.array([[0, 1, 2], [0, 2, 3]], np.uint32) + 81
    verts, faces = np.concatenate([gv, rv]), np.concatenate([gf, rf])
    pts = np.stack([X.ravel() * 0.75 + 1.0, Y.ravel() * 0.75 + 1.0, np.zeros(81, np.float32)], 1).astype(np.float32)
    nrm = np.tile(np.array([0, 0, 1], np.float32), (81, 1))
    specs = [((float(p[0]), float(p[1]), 20.0), 0.0, 0.0) for p in pts[::4]]      # exactly above every 4th point
    pos, w2c, calib = scenes._pack_views(specs)
    sc = scenes.Scene("flat", verts, faces, pts, nrm, np.zeros(81, np.float32), np.ones(81, np.float32), pos, w2c, calib)
    for arrays in (oracle.BVH.build(verts, faces, threads=1).export(),):
        obs_o, data, rows, state = _populate_both(api, oracle, sc, arrays)
        assert (state == 1).sum() > 0 and (state == 2).sum() > 0
        assert np.array_equal(rows, obs_o.num_rows)
        assert _valid_rows_equal(obs_o.data, data, rows, 32)
    gb = api.BVHTree.build_lbvh(verts, faces)
    nodes, aabbs, tris, indices = gb.export()
    arrays = oracle.TreeArrays(nodes, aabbs, tris, indices)
    obs_o, data, rows, state = _populate_both(api, oracle, sc, arrays)
    assert np.array_equal(rows, obs_o.num_rows)


@pytest.mark.parametrize("early_at,split_at", [("1", "16"), ("8", "64"), ("0", "1024")])
def test_cull_early_list_and_view_range_pieces_are_bit_exact(api, oracle, early_at, split_at, monkeypatch):
    """k_cull_emit_q starts the blocks with many candidate views first and cuts the largest into up to 8 view-range pieces
    whose records the last piece moves together.  With the thresholds lowered every block of a small scene with 700 views
    takes that path (pieces of 128 views, some of them empty); rows, counters and order must not change."""
    from uavmvs_b200 import scenes
    sc = scenes.urban(num_tris=8000, num_points=8192 + 17, num_views=700, extent=64.0, seed=6)
    monkeypatch.setenv("UAVMVS_CULL_EARLY_AT", early_at)
    monkeypatch.setenv("UAVMVS_CULL_SPLIT_AT", split_at)
    gbvh = api.BVHTree.build_sah(sc.verts, sc.faces)
    nodes, aabbs, tris, indices = gbvh.export()
    obvh = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    obs_o, state, _ = _run_oracle_eval(oracle, obvh, sc)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, sc.views28(), sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    data, rows = obs.download()
    assert rows.max() > 32
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)
    # removal goes through the same kernel
    api.update_observation_rays_batch(False, sc.views28()[100:400], sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    oracle.update_observation_rays(obvh, False, sc.views28()[100:400], sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    api.process_observation_rays(obs)
    data, rows = obs.download()
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)


@pytest.mark.parametrize("tau", ["4", "32"])
def test_packet_entry_points_opt_in_is_bit_exact(api, oracle, tau, monkeypatch):
    """UAVMVS_ENTRY=1: rays start from their packet's frontier of wide nodes (k_entry) instead of the root."""
    from uavmvs_b200 import scenes
    sc = scenes.dense()
    monkeypatch.setenv("UAVMVS_ENTRY", "1")
    monkeypatch.setenv("UAVMVS_ENTRY_TAU", tau)
    gbvh = api.BVHTree.build_ploc(sc.verts, sc.faces)
    nodes, aabbs, tris, indices = gbvh.export()
    obvh = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    obs_o, state, _ = _run_oracle_eval(oracle, obvh, sc)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, sc.views28(), sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    data, rows = obs.download()
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)


@pytest.mark.parametrize("scene_name", ["tiny", "dense"])
def test_visibility_shafts_opt_in_is_bit_exact(api, oracle, scene_name, monkeypatch):
    """UAVMVS_SHAFTS=1: (block, view) packets certified visible by the shaft pass skip the per-ray
    trace; counts and stored rows must stay bit-identical, and some packets must be certified."""
    from uavmvs_b200 import scenes
    sc = getattr(scenes, scene_name)()
    monkeypatch.setenv("UAVMVS_SHAFTS", "1")
    gbvh = api.BVHTree.build_ploc(sc.verts, sc.faces)
    nodes, aabbs, tris, indices = gbvh.export()
    obvh = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    obs_o, state, _ = _run_oracle_eval(oracle, obvh, sc)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    api.k1_counters(reset=True)
    api.update_observation_rays_batch(True, sc.views28(), sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    data, rows = obs.download()
    packets, certified, rays_cert, rays_traced = api.k1_counters()
    assert packets > 0 and rays_cert + rays_traced == int((state != 0).sum()), (packets, certified, rays_cert, rays_traced)
    if scene_name == "dense":
        assert certified > 0 and rays_cert > 0, (packets, certified, rays_cert, rays_traced)
    assert np.array_equal(rows, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data, rows, 32)
    # and the removal pass (populate = false) through the same packets
    api.update_observation_rays_batch(False, sc.views28()[::3], sc.max_distance, sc.width, sc.height, gbvh, cloud, obs)
    oracle.update_observation_rays(obvh, False, sc.views28()[::3], sc.max_distance, sc.width, sc.height,
                                   sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    api.process_observation_rays(obs)
    data2, rows2 = obs.download()
    assert np.array_equal(rows2, obs_o.num_rows)
    assert _valid_rows_equal(obs_o.data, data2, rows2, 32)


@pytest.mark.parametrize("builder", ["sah", "build_ploc"])
def test_closest_point_equals_oracle(api, oracle, tiny_scene, tiny_tree, builder):
    """uavmvs_bvh_closest_point vs the oracle's acc::BVHTree::closest_point (pinned to the reference's own code
    in tests/test_oracle_vs_ref.py) on the same tree: found flag, leaf-order triangle index, point bit for bit."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_golden_closest as mc
    sc = tiny_scene
    if builder == "sah":
        obvh, arrays = tiny_tree
        gbvh = _gpu_tree(api, arrays)
    else:
        gbvh = api.BVHTree.build_ploc(sc.verts, sc.faces)
        nodes, aabbs, tris, indices = gbvh.export()
        obvh = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    q = mc.queries(sc)
    for k, md in enumerate(mc.MAX_DISTS):
        f_o, i_o, c_o = obvh.closest_point(q, md)
        f_g, i_g, c_g = gbvh.closest_point(q, md)
        assert np.array_equal(f_g, f_o) and np.array_equal(i_g, i_o)
        assert np.array_equal(c_g.view(np.uint32), c_o.view(np.uint32))
        if builder == "sah":                                       # and the committed reference output itself
            g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_acc_closest.npz"))
            assert np.array_equal(np.packbits(f_g), g["found%d" % k]) and np.array_equal(i_g, g["idx%d" % k])
            assert np.array_equal(c_g.view(np.uint32), g["cp%d" % k].view(np.uint32))
    # single-triangle tree (the root is a leaf) and an empty query list
    v = np.array([[0, 0, 0], [2, 0, 0], [0, 2, 0]], np.float32)
    one = oracle.BVH.build(v, np.array([[0, 1, 2]], np.uint32), threads=1)
    g1 = _gpu_tree(api, one.export())
    pts = np.array([[0.5, 0.5, 3.0], [5.0, 5.0, 0.0], [-1.0, -1.0, -1.0]], np.float32)
    a, b = one.closest_point(pts), g1.closest_point(pts)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2].view(np.uint32), b[2].view(np.uint32))
    assert g1.closest_point(np.zeros((0, 3), np.float32))[0].shape == (0,)


def test_config1_bundled_mesh_full_evaluation(api, oracle, tmp_path):
    """BASELINE.json configs[0]: the reference's bundled res/meshes/sphere.ply (x30, on a ground quad), 100k
    proxy points, 200 views, evaluated on the tree of the reference's CPU path (acc SAH builder, pinned to the
    reference's own code) - driven through the FILES a uavmvs user has (PLY mesh, PLY cloud, trajectory text)."""
    from uavmvs_b200 import scenes, formats as F
    sc = scenes.config1()
    assert sc.num_tris == 20482 and sc.num_points == 100_000 and 150 <= sc.num_views <= 250
    F.save_ply(str(tmp_path / "mesh.ply"), sc.verts, sc.faces)
    F.save_ply(str(tmp_path / "cloud.ply"), sc.points, None, sc.normals)
    F.save_trajectory(str(tmp_path / "t.traj"), sc.view_pos, sc.w2c[:, :3, :3], np.full(sc.num_views, scenes.FLEN, np.float32))
    mesh = F.load_ply(str(tmp_path / "mesh.ply"))
    pts, nrm, val, qual = F.load_point_cloud(str(tmp_path / "cloud.ply"))
    views = F.views_from_trajectory(*F.load_trajectory(str(tmp_path / "t.traj")), sc.width, sc.height)
    obvh = oracle.BVH.build(mesh["vertices"], mesh["faces"], threads=4)          # libs/acc/bvh_tree.h:335-399
    gbvh = _gpu_tree(api, obvh.export())
    cloud = api.PointCloud.create(pts, nrm, val, qual)
    obs = api.VectorArray.create(len(pts), 32)
    recons, wrecons = api.Array.create(len(pts)), api.Array.create(len(pts))
    s_sum, s_min, s_max, stats = api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height,
                                                         gbvh, cloud, obs, recons, wrecons)
    obs_o = oracle.ObsRays(len(pts), 32)
    _, cnt = oracle.update_observation_rays(obvh, True, views, sc.max_distance, sc.width, sc.height, pts, nrm, obs_o,
                                            want_counters=True)
    oracle.process_observation_rays(obs_o)
    r_o = oracle.evaluate_observation_rays(obs_o)
    data, rows = obs.download()
    assert stats[0] == int(cnt[0]) > 300_000
    assert np.array_equal(rows, obs_o.num_rows) and _valid_rows_equal(obs_o.data, data, rows, 32)
    # recon: 1e-5 relative (north_star) plus half an ulp of 1.0 per pair: matchability is (1 - logistic), which
    # cancels for wide-parallax pairs, so a point whose ONLY pair is such a pair (recon ~ 1e-5..1e-4 here) carries
    # the rounding of the logistic itself (libm expf / IEEE division in the oracle, __expf in the reference and
    # here) at ~1e-4 relative.  11 of the 100 000 points are in that regime; everything else passes 1e-5 alone.
    r_g = recons.download().astype(np.float64)
    n = np.minimum(rows, 32).astype(np.float64)
    tol = 1e-5 * np.maximum(np.abs(r_g), np.abs(r_o)) + 3e-8 * n * (n - 1) / 2
    assert np.all(np.abs(r_g - r_o) <= tol)
    strict = np.abs(r_g - r_o) <= 1e-5 * np.maximum(np.abs(r_g), np.abs(r_o))
    assert strict.mean() > 0.9995
    w_o = oracle.calculate_func_recons(r_o, sc.target_recon).astype(np.float64)
    assert abs(s_sum / len(pts) - w_o.mean()) <= 1e-5 * w_o.mean() + 1e-9


def test_concurrent_populate_on_one_observation_array(api, oracle):
    """The reference's optimiser calls update_observation_rays(true, ...) for different views from several host threads /
    streams on the SAME obs_rays (optimize_trajectory.cu:470-495); rows are taken with atomicAdd there.  Two host threads,
    two streams, the two halves of a dense trajectory (views that share points and 32-point blocks): counters equal the
    serial result, and so does the multiset of rows of every column that never overflowed."""
    import threading
    import torch
    from uavmvs_b200 import scenes
    sc = scenes.dense()
    bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    views = sc.views28()
    serial = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, views, sc.max_distance, sc.width, sc.height, bvh, cloud, serial)
    d_s, r_s = serial.download()
    for attempt in range(4):
        obs = api.VectorArray.create(sc.num_points, 32)
        obs.null()
        api.synchronize()
        errors = []

        def work(part):
            try:
                api.set_device(0)
                st = torch.cuda.Stream()
                for k in range(0, part.shape[0], 8):         # several launches per thread: more interleavings
                    api.update_observation_rays_batch(True, part[k:k + 8], sc.max_distance, sc.width, sc.height, bvh, cloud,
                                                      obs, stream=st)
                api.synchronize(st)
            except Exception as e:                           # noqa: BLE001
                errors.append(e)
        th = [threading.Thread(target=work, args=(views[0::2],)), threading.Thread(target=work, args=(views[1::2],))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert not errors, errors
        d_c, r_c = obs.download()
        assert np.array_equal(r_c, r_s)                      # no lost increment, no row handed out twice
        ok = r_s <= 32
        m = (np.arange(32)[:, None] < np.minimum(r_s, 32)[None, :]) & ok[None, :]

        def keys(d):
            a = np.where(m[:, :, None], d, 0).view(np.uint32).astype(np.uint64)
            return np.sort(a[:, :, 0] * 7 + a[:, :, 1] * 13 + a[:, :, 2] * 31 + a[:, :, 3], axis=0)
        assert np.array_equal(keys(d_c), keys(d_s))


def test_tree_deeper_than_the_stacks_is_refused(api):
    """A chain-shaped tree of 240 levels (valid indices, nested boxes) must come back as an error from the upload, not
    overflow the traversal stacks (the reference has the same unchecked stacks, libs/cacc/tracing.h:28-33)."""
    NAI = 0xFFFFFFFF
    depth = 240
    tris = np.zeros((depth + 1, 9), np.float32)
    for t in range(depth + 1):
        tris[t] = [t, 0, 0, t + 0.5, 0, 0, t, 0.5, 0]
    nodes, aabbs = [], []
    # node 2k: inner {leaf 2k+1, inner 2k+2}; the last inner node holds two leaves
    for k in range(depth):
        last = k == depth - 1
        nodes.append([k, depth + 1, 2 * k + 1, 2 * k + 2])
        aabbs.append([k, 0, 0, depth + 0.5, 0.5, 0])
        nodes.append([k, k + 1, NAI, NAI])
        aabbs.append([k, 0, 0, k + 0.5, 0.5, 0])
        if last:
            nodes.append([k + 1, k + 2, NAI, NAI])
            aabbs.append([k + 1, 0, 0, k + 1.5, 0.5, 0])
    with pytest.raises(api.UavmvsError, match="deeper than the traversal stacks"):
        api.BVHTree.create(np.array(nodes, np.uint32), np.array(aabbs, np.float32), tris, np.arange(depth + 1, dtype=np.uint32))


def test_heuristic_constants_per_observation_array(api, oracle, tiny_scene, tiny_tree):
    """uavmvs_obs_configure_heuristic: two observation arrays evaluated with different constants in one process
    (per context, not global); each equals the oracle run with its own constants."""
    sc = tiny_scene
    obvh, arrays = tiny_tree
    gbvh = _gpu_tree(api, arrays)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs_o = oracle.ObsRays(sc.num_points, 32)
    oracle.update_observation_rays(obvh, True, sc.views28(), sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    oracle.process_observation_rays(obs_o)
    heurs = [(8.0, 4.0, 32.0, 16.0), (4.0, 3.0, 16.0, 12.0)]
    arrays_g = []
    for h in heurs:
        obs = api.VectorArray.create(sc.num_points, 32)
        obs.configure_heuristic(*h)
        arrays_g.append(obs)
    recons = [api.Array.create(sc.num_points) for _ in heurs]
    wrecons = [api.Array.create(sc.num_points) for _ in heurs]
    for obs, rc, wr in zip(arrays_g, recons, wrecons):      # interleaved use of the two contexts
        api.evaluate_trajectory(sc.views28(), sc.max_distance, sc.target_recon, sc.width, sc.height, gbvh, cloud, obs, rc, wr)
    got = [rc.download() for rc in recons]
    for h, r_g in zip(heurs, got):
        r_o = oracle.evaluate_observation_rays(obs_o, heur=h)
        assert np.all(np.abs(r_g.astype(np.float64) - r_o) <= 1e-5 * np.maximum(np.abs(r_g), np.abs(r_o)) + 1e-6)
    assert not np.allclose(got[0], got[1])
