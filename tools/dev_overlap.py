"""Dev experiment: does running the evaluation of K point shards on K host threads / streams
concurrently beat one evaluation of the whole cloud?  (overlap of issue-bound and latency-bound kernels)"""
import argparse, os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from uavmvs_b200 import api, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--shards", type=int, default=2)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--interleave", type=int, default=4096)
args = ap.parse_args()
sc = scenes.urban(1_000_000, 1_000_000, 2000, 256.0)
torch.cuda.init()
bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
views = api.DeviceViews(sc.views28())
K = args.shards
n = sc.num_points
chunk = np.arange(n) // args.interleave
shards = []
for k in range(K):
    idx = np.nonzero(chunk % K == k)[0]
    pts, nrm = np.ascontiguousarray(sc.points[idx]), np.ascontiguousarray(sc.normals[idx])
    shards.append(dict(cloud=api.PointCloud.create(pts, nrm), obs=api.VectorArray.create(len(idx), 32),
                       recons=api.Array.create(len(idx)), wrecons=api.Array.create(len(idx)),
                       stream=torch.cuda.Stream(), n=len(idx)))
out = [None] * K
start, done = threading.Barrier(K + 1), threading.Barrier(K + 1)
def worker(i):
    sh = shards[i]
    for it in range(args.iters):
        start.wait()
        out[i] = api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height, bvh, sh["cloud"],
                                         sh["obs"], sh["recons"], sh["wrecons"], stream=sh["stream"])
        api.synchronize(sh["stream"])
        done.wait()
th = [threading.Thread(target=worker, args=(i,)) for i in range(K)]
for t in th: t.start()
for it in range(args.iters):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    start.wait()
    done.wait()
    ms = (time.perf_counter() - t0) * 1e3
    rays = sum(o[3][0] for o in out)
    print("shards %d iter %d: %.2f ms wall  rays %d -> %.3f Grays/s score %.5f" % (K, it, ms, rays, rays / ms / 1e6,
          sum(o[0] for o in out) / n), flush=True)
for t in th: t.join()
