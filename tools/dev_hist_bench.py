"""Dev timing of the spherical-histogram path on the config-2 scene (configs 3 / 4 of BASELINE.json):
batched K4 + K6 over B candidate positions vs the per-position entry points (the reference's loop)."""
import argparse, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from uavmvs_b200 import api, scenes
import oracle

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=4096)
ap.add_argument("--single", type=int, default=64)
args = ap.parse_args()
sc = scenes.urban(1_000_000, 1_000_000, 2000, 256.0)
torch.cuda.init()
bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
rng = np.random.default_rng(4)
values = rng.uniform(0.0, 1.2, sc.num_points).astype(np.float32)
quals = rng.uniform(0.0, 1.0, sc.num_points).astype(np.float32)
cloud = api.PointCloud.create(sc.points, sc.normals, values, quals)
sv = oracle.sphere_vertices(1.0, 3)
kn, kv = oracle.KD.build(sv).export()
kd = api.KDTree.create(kn, kv)
S = sv.shape[0]
B = args.batch
pos = rng.uniform([0, 0, 30], [256, 256, 90], (B, 3)).astype(np.float32)
hists = api.Array.create(B * S)
def ev():
    return torch.cuda.Event(enable_timing=True)
for it in range(3):
    hists.null()
    torch.cuda.synchronize()
    e0, e1, e2 = ev(), ev(), ev()
    e0.record()
    api.populate_spherical_histogram_batch(pos, sc.max_distance, bvh, cloud, kd, hists)
    e1.record()
    best = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, kd, hists, B)
    e2.record()
    torch.cuda.synchronize()
    c = api.k1_counters()
    print("batched: %d positions  K4 %.2f ms (%.1f us/position, %.2f Grays/s, %d rays)  K6+argmin %.2f ms (%.2f us/position)" % (
        B, e0.elapsed_time(e1), 1e3 * e0.elapsed_time(e1) / B, c[3] / e0.elapsed_time(e1) / 1e6, c[3], e1.elapsed_time(e2),
        1e3 * e1.elapsed_time(e2) / B), flush=True)
h1 = api.Array.create(S)
img = api.Image.create(128, 45)
n = args.single
torch.cuda.synchronize()
e0, e1 = ev(), ev()
e0.record()
for b in range(n):
    h1.null()
    api.populate_spherical_histogram(pos[b], sc.max_distance, bvh, cloud, kd, h1)
    api.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, kd, h1, img)
e1.record()
torch.cuda.synchronize()
print("per-position entry points (reference launch pattern): %.1f us/position over %d positions" % (1e3 * e0.elapsed_time(e1) / n, n))
