"""Dev timing of K7 estimate_capture_difficulty (closest-hit hemisphere rays) on the config-2 scene."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from uavmvs_b200 import api, scenes
import oracle

ap = argparse.ArgumentParser()
ap.add_argument("--points", type=int, default=100_000)
ap.add_argument("--check", type=int, default=2000)
args = ap.parse_args()
sc = scenes.urban(1_000_000, 1_000_000, 16, 256.0)
torch.cuda.init()
bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
n = args.points
idx = np.sort(np.random.default_rng(0).choice(sc.num_points, n, replace=False))
pts, nrm = sc.points[idx], sc.normals[idx]
sv = oracle.sphere_vertices(1.0, 3)
kn, kv = oracle.KD.build(sv).export()
kd = api.KDTree.create(kn, kv)
cloud = api.PointCloud.create(pts, nrm)
dirs = sv / np.linalg.norm(sv, axis=1, keepdims=True)
rays = int(((nrm @ dirs.T) >= 0.087).sum())
for it in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    api.estimate_capture_difficulty(sc.max_distance, bvh, sc.num_tris, kd, cloud)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print("K7: %d points, %d closest-hit rays, %.2f ms -> %.3f Grays/s" % (n, rays, ms, rays / ms / 1e6), flush=True)
if args.check:
    v_g, q_g = cloud.download_attributes()
    nodes, aabbs, tris, indices = bvh.export()
    ob = oracle.BVH.from_arrays(oracle.TreeArrays(nodes, aabbs, tris, indices))
    m = args.check
    v_o, q_o = oracle.estimate_capture_difficulty(ob, sc.max_distance, sc.num_tris, sv, pts[:m], nrm[:m])
    print("check %d points: values max abs diff %.3g, qualities max abs diff %.3g" % (
        m, np.abs(v_g[:m] - v_o).max(), np.abs(q_g[:m] - q_o).max()))
