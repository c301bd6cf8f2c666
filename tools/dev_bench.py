"""Developer timing harness (not the judged bench): config-2-like scene, GPU LBVH,
evaluate_trajectory timed with CUDA events."""
import argparse
import sys
import time
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from uavmvs_b200 import api, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--tris", type=int, default=1_000_000)
ap.add_argument("--points", type=int, default=1_000_000)
ap.add_argument("--views", type=int, default=2000)
ap.add_argument("--extent", type=float, default=256.0)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--check", action="store_true")
ap.add_argument("--ploc", action="store_true")
ap.add_argument("--gsah", action="store_true", help="GPU sweep-SAH builder")
ap.add_argument("--sah", action="store_true", help="dev experiment: tree from the oracle SAH builder (test infrastructure)")
args = ap.parse_args()

t = time.time()
sc = scenes.urban(args.tris, args.points, args.views, args.extent)
print("scene", sc.name, "gen %.1fs" % (time.time() - t), flush=True)
torch.cuda.init()
t = time.time()
if args.sah:
    import oracle
    ta = oracle.BVH.build(sc.verts, sc.faces).export()
    bvh = api.BVHTree.create(ta.nodes, ta.aabbs, ta.tris, ta.indices)
elif args.gsah:
    bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
elif args.ploc:
    bvh = api.BVHTree.build_ploc(sc.verts, sc.faces)
else:
    bvh = api.BVHTree.build_lbvh(sc.verts, sc.faces)
api.synchronize()
print("lbvh build+upload %.3fs" % (time.time() - t), bvh.sizes(), flush=True)
cloud = api.PointCloud.create(sc.points, sc.normals)
obs = api.VectorArray.create(sc.num_points, 32)
recons, wrecons = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
dviews = api.DeviceViews(sc.views28())
stream = torch.cuda.current_stream()
for it in range(args.iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); t_w = time.perf_counter()
    e0.record(stream)
    s, mn, mx, st = api.evaluate_trajectory(dviews, sc.max_distance, sc.target_recon, sc.width, sc.height,
                                            bvh, cloud, obs, recons, wrecons, stream=stream)
    e1.record(stream)
    torch.cuda.synchronize()
    print("wall %.2f ms" % ((time.perf_counter() - t_w) * 1e3))
    ms = e0.elapsed_time(e1)
    print("stages", ["%.3f" % x for x in api.profile_last()], "k1 counters [packets, certified, rays certified, rays traced]", api.k1_counters())
    print("iter %d: %.2f ms  rays %d vis %d pairs %d after_block_cull %d  -> %.3f Grays/s %.2f Gpairs/s score %.5f" % (
        it, ms, st[0], st[1], st[2], st[3], st[0] / ms / 1e6, st[2] / ms / 1e6, s / sc.num_points), flush=True)
rows = obs.download(data=False)[1]
print("rows mean %.2f max %d frac>=32 %.3f" % (rows.mean(), rows.max(), (rows >= 32).mean()))
if args.check:
    import oracle
    nodes, aabbs, tris, idx = bvh.export()
    ob = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, idx))
    sub = min(sc.num_views, 64)
    o = oracle.ObsRays(sc.num_points)
    t = time.time()
    _, cnt = oracle.update_observation_rays(ob, True, sc.views28()[:sub], sc.max_distance, sc.width, sc.height,
                                            sc.points, sc.normals, o, want_counters=True)
    dt = time.time() - t
    print("oracle %d views: %.1fs rays %d inner/ray %.1f leaf/ray %.2f tri/ray %.2f -> %.3f Mrays/s" % (
        sub, dt, cnt[0], cnt[1] / cnt[0], cnt[2] / cnt[0], cnt[3] / cnt[0], cnt[0] / dt / 1e6))
    obs2 = api.VectorArray.create(sc.num_points, 32)
    api.update_observation_rays_batch(True, sc.views28()[:sub], sc.max_distance, sc.width, sc.height, bvh, cloud, obs2)
    r2 = obs2.download(data=False)[1]
    print("counts equal on subsample:", np.array_equal(r2, o.num_rows), int(r2.sum()))
