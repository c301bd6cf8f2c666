"""Point-sharded evaluation over the GPUs of one box (one process per GPU).

Every output of the path is per proxy point and depends only on that point plus
replicated read-only data (BVH, KD-tree, trajectory), so the proxy cloud is cut into
contiguous Z-order ranges, one per rank, and there is no collective on the data path
(SURVEY.md 8e).  The only exchanges are

  * the trajectory score: ONE all-gather of an 8-double record per rank {sum wrecon, #points, rays,
    observations, pairs, min, max}, folded in rank order -- what cacc::reduction::sum/min/max
    (libs/cacc/reduction.h:132-162) produce on one GPU in the reference.  In the product this is
    uavmvs_evaluate_trajectory_sharded (C ABI, ncclAllGather on the evaluation's stream);
  * on request the full per-point vector (recon, counts): an all-gather of the shards;
  * for the histogram kernels one SUM all-reduce of the [B, S] partial histograms.

torch.distributed is the plumbing (NCCL over NVLink on the GPU box, gloo in the CPU
tests); nothing here computes on the host.
"""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous range [lo, hi) of a Z-ordered cloud of n points owned by `rank`."""
    per = (n + world - 1) // world
    return min(rank * per, n), min((rank + 1) * per, n)


def _dist():
    import torch.distributed as dist
    return dist


def _device_for(group=None):
    import torch
    dist = _dist()
    if dist.is_initialized() and dist.get_backend(group) == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def fold_records(recs):
    """Fold of the per-rank score records {sum, points, rays, observations, pairs after the block cull, min, max,
    pairs} in rank order -- the host mirror of k_fold_score (csrc/eval.cu)."""
    tot = pts = rays = obs = cull = pairs = 0.0
    mn, mx = float("inf"), float("-inf")
    for q in recs:
        tot += q[0]; pts += q[1]; rays += q[2]; obs += q[3]; cull += q[4]; pairs += q[7]
        mn, mx = min(mn, q[5]), max(mx, q[6])
    pts = int(round(pts))
    return {"sum": tot, "min": mn, "max": mx, "points": pts, "score": tot / max(pts, 1),
            "stats": [int(round(rays)), int(round(obs)), int(round(pairs)), int(round(cull))]}


def combine_scores(local_sum, local_min, local_max, local_points, stats=None, group=None):
    """Per-shard reduction of wrecon -> trajectory-level result with ONE collective: an all-gather of the 8-double
    record of every rank, folded in rank order (what uavmvs_evaluate_trajectory_sharded does on the device with
    ncclAllGather; this torch.distributed form serves host-side callers and the gloo tests).

    Returns dict(sum, min, max, points, score, stats) with score = sum / points
    (apps/optimize_trajectory/optimize_trajectory.cu:529).  Without an initialised
    process group this is the single-GPU identity."""
    import torch
    stats = list(stats) if stats is not None else [0, 0, 0, 0]
    rec = [float(local_sum), float(local_points), float(stats[0]), float(stats[1]), float(stats[3]),
           float(local_min), float(local_max), float(stats[2])]
    dist = _dist()
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return fold_records([rec])
    dev = _device_for(group)
    mine = torch.tensor(rec, dtype=torch.float64, device=dev)
    every = torch.empty(dist.get_world_size(group) * 8, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(every, mine, group=group)
    return fold_records(every.cpu().reshape(-1, 8).tolist())


def gather_point_vector(local, n_total, group=None):
    """All-gather a per-point vector (recon, counts) from the shards, in cloud order."""
    import torch
    dist = _dist()
    local = np.ascontiguousarray(local)
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return local.copy()
    if local.dtype == np.uint32:      # not a collective dtype everywhere
        return gather_point_vector(local.astype(np.int64), n_total, group).astype(np.uint32)
    world = dist.get_world_size(group)
    per = (n_total + world - 1) // world
    dev = _device_for(group)
    pad = np.zeros(per, local.dtype)
    pad[:local.shape[0]] = local
    t = torch.from_numpy(pad).to(dev)
    out = torch.empty(world * per, dtype=t.dtype, device=dev)
    dist.all_gather_into_tensor(out, t, group=group)
    return out.cpu().numpy()[:n_total]


def allreduce_histograms(hist, group=None):
    """SUM all-reduce of partial sphere histograms [B, S] (torch tensor, in place)."""
    dist = _dist()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    return hist


class ShardedEvaluator:
    """evaluate_trajectory over this rank's shard of the proxy cloud.

    The mesh is replicated: every rank builds (or uploads) the same BVH; `points` /
    `normals` are the FULL cloud, of which this rank keeps [lo, hi)."""

    def __init__(self, api, bvh_tree, points, normals, rank=0, world=1, max_rows=32, values=None, qualities=None):
        self.api, self.bvh, self.rank, self.world = api, bvh_tree, rank, world
        self.n_total = int(np.asarray(points).reshape(-1, 3).shape[0])
        self.lo, self.hi = shard_range(self.n_total, rank, world)
        sl = slice(self.lo, self.hi)
        pts = np.asarray(points, np.float32).reshape(-1, 3)[sl]
        nrm = np.asarray(normals, np.float32).reshape(-1, 3)[sl]
        self.n_local = pts.shape[0]
        self.cloud = api.PointCloud.create(pts, nrm,
                                           None if values is None else np.asarray(values, np.float32)[sl],
                                           None if qualities is None else np.asarray(qualities, np.float32)[sl])
        self.obs = api.VectorArray.create(self.n_local, max_rows)
        self.recons, self.wrecons = api.Array.create(self.n_local), api.Array.create(self.n_local)

    def evaluate_local(self, views, max_distance, target_recon, width, height, stream=None):
        return self.api.evaluate_trajectory(views, max_distance, target_recon, width, height, self.bvh, self.cloud,
                                            self.obs, self.recons, self.wrecons, stream=stream)

    def evaluate(self, views, max_distance, target_recon, width, height, stream=None, comm=None, group=None):
        """Shard evaluation + the score collective.  With `comm` (api.Comm, or None on one GPU) the whole thing is
        one C-ABI call (uavmvs_evaluate_trajectory_sharded, ncclAllGather on the stream); with a torch.distributed
        `group` and no comm the collective goes through torch.distributed instead."""
        if group is None:
            res, _ = self.api.evaluate_trajectory_sharded(comm, views, max_distance, target_recon, width, height,
                                                          self.bvh, self.cloud, self.obs, self.recons, self.wrecons,
                                                          stream=stream)
            res["stats"] = [res["rays"], res["observations"], res["pairs"], res["pairs_after_block_cull"]]
            return res
        s, mn, mx, st = self.evaluate_local(views, max_distance, target_recon, width, height, stream)
        return combine_scores(s, mn, mx, self.n_local, st, group)

    def gather_recons(self, group=None):
        return gather_point_vector(self.recons.download(), self.n_total, group)

    def gather_counts(self, group=None):
        return gather_point_vector(self.obs.download(data=False)[1], self.n_total, group)
