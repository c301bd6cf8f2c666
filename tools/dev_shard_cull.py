"""Dev: config 5's strong-scaling shards of ranks r of `world`, run ONE AFTER THE OTHER on a single GPU with the per-kernel
K1 profile, to look at per-rank imbalance without an N-GPU box.  Not part of any bench line.

    python tools/dev_shard_cull.py --world 8 --ranks 0,1,2 [--debug]
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--ranks", default="0,1")
    ap.add_argument("--debug", action="store_true", help="one more evaluation with UAVMVS_K1_DEBUG=1 (per-block cull cycles)")
    ap.add_argument("--reps", type=int, default=4)
    ap.add_argument("--chunk", type=int, default=0, help="points per interleaved chunk (0: bench.py's CHUNK)")
    a = ap.parse_args()
    if a.chunk:
        bench.CHUNK = a.chunk
    from uavmvs_b200 import api, scenes
    sc = scenes.urban(10_000_000, 8_000_000, 10_000, 800.0, seed=1)
    t0 = time.time()
    bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
    print("tree %.2f s" % (time.time() - t0), flush=True)
    views = api.DeviceViews(sc.views28())

    class A:
        scaling = "strong"
        shard_chunk = bench.CHUNK if bench.CHUNK < 1000000 else 0      # >= 1M: contiguous ranges
    for r in [int(x) for x in a.ranks.split(",")]:
        p, n = bench.rank_cloud(A, sc, r, a.world)
        cloud = api.PointCloud.create(p, n)
        obs = api.VectorArray.create(p.shape[0], 32)
        rc, wr = api.Array.create(p.shape[0]), api.Array.create(p.shape[0])
        api.profile_k1_enable(True)
        best = None
        for i in range(a.reps):
            obs.null()
            api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height, bvh, cloud, obs, rc, wr)
            ms = api.profile_k1()
            if i and (best is None or sum(ms) < sum(best)):
                best = ms
        print("chunk %d " % bench.CHUNK, end="")
        print("rank %d of %d: %d points  masks %.3f caps %.3f cull %.3f trace %.3f retire %.3f  K1 %.3f" %
              (r, a.world, p.shape[0], *best, sum(best)), flush=True)
        if a.debug:
            os.environ["UAVMVS_K1_DEBUG"] = "1"
            obs.null()
            api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height, bvh, cloud, obs, rc, wr)
            api.k1_counters(True)
            del os.environ["UAVMVS_K1_DEBUG"]
        del cloud, obs, rc, wr


if __name__ == "__main__":
    main()
