"""Multi-GPU entry of the C ABI (uavmvs_evaluate_trajectory_sharded): the shard evaluation followed by ONE NCCL
collective on the same stream, through an ncclComm_t created by the library itself.

  world 1  runs on any GPU box: a one-rank NCCL communicator still goes through ncclAllGather, so the dlopen'ed
           NCCL, the packed record and the device fold are exercised by the driver's single-GPU `pytest -m gpu`.
  world 2  one process per GPU (skipped on a one-GPU box): two point shards of the tiny scene; both ranks must
           return the same bits, equal to the fold of the two single-GPU evaluations.
"""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _evaluate_unsharded(api, sc, bvh, pts, nrm):
    cloud = api.PointCloud.create(pts, nrm)
    obs = api.VectorArray.create(pts.shape[0], 32)
    rc, wr = api.Array.create(pts.shape[0]), api.Array.create(pts.shape[0])
    s, mn, mx, st = api.evaluate_trajectory(sc.views28(), sc.max_distance, sc.target_recon, sc.width, sc.height, bvh,
                                            cloud, obs, rc, wr)
    return dict(sum=s, min=mn, max=mx, stats=st, recons=rc.download(), rows=obs.download(data=False)[1],
                wsum64=float(wr.download().astype(np.float64).sum()))


def test_sharded_entry_world1_nccl(tiny_scene):
    import torch  # noqa: F401  (loads the NCCL the library then finds with dlopen)
    from uavmvs_b200 import api
    api.set_device(0)
    sc = tiny_scene
    bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
    ref = _evaluate_unsharded(api, sc, bvh, sc.points, sc.normals)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    rc, wr = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
    h_rc, h_rows = np.empty(sc.num_points, np.float32), np.empty(sc.num_points, np.uint32)
    comm = api.Comm.create(api.Comm.unique_id(), 1, 0)
    for c in (None, comm):                                  # without NCCL, then through a one-rank communicator
        for views in (sc.views28(), api.DeviceViews(sc.views28())):
            score, loc = api.evaluate_trajectory_sharded(c, views, sc.max_distance, sc.target_recon, sc.width, sc.height,
                                                         bvh, cloud, obs, rc, wr, h_recons=h_rc, h_num_rows=h_rows)
            assert score["points"] == sc.num_points and score["rays"] == ref["stats"][0]
            assert score["observations"] == ref["stats"][1] and score["pairs"] == sc.num_points * sc.num_views
            assert score["pairs_after_block_cull"] == ref["stats"][3]
            assert np.float32(score["sum"]) == np.float32(ref["sum"]) and score["min"] == ref["min"] and score["max"] == ref["max"]
            assert abs(score["sum"] - ref["wsum64"]) <= 1e-6 * abs(ref["wsum64"])      # K9: float partial sums, double across blocks
            assert loc[0] == ref["sum"]
            assert np.array_equal(h_rc.view(np.uint32), ref["recons"].view(np.uint32)) and np.array_equal(h_rows, ref["rows"])
    assert api.profile_last_collective() >= 0.0
    # histogram all-reduce and per-point all-gather degenerate to identities on one rank
    a = api.Array.create(64)
    a.upload(np.arange(64, dtype=np.float32))
    api.allreduce_array(comm, a, 64)
    out = api.Array.create(64)
    api.allgather_array(comm, a, 64, out)
    api.synchronize()
    assert np.array_equal(out.download(), np.arange(64, dtype=np.float32))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from uavmvs_b200 import api, scenes, sharded
    torch.cuda.set_device(rank)
    api.set_device(rank)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    comm = api.Comm.from_torch_distributed()
    sc = scenes.tiny()
    bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
    ev = sharded.ShardedEvaluator(api, bvh, sc.points, sc.normals, rank, world)
    res = ev.evaluate(sc.views28(), sc.max_distance, sc.target_recon, sc.width, sc.height, comm=comm)
    # partial histograms: SUM all-reduce through the C ABI
    h = api.Array.create(16)
    h.upload(np.full(16, float(rank + 1), np.float32))
    api.allreduce_array(comm, h, 16)
    # per-point vector: all-gather in shard order (shards padded to the same length)
    per = (sc.num_points + world - 1) // world
    pad = api.Array.create(per)
    r = ev.recons.download()
    padded = np.zeros(per, np.float32)
    padded[:r.shape[0]] = r
    pad.upload(padded)
    allv = api.Array.create(per * world)
    api.allgather_array(comm, pad, per, allv)
    api.synchronize()
    np.savez(os.path.join(out_dir, "rank%d.npz" % rank), rec=np.array([res[k] for k in api.SCORE_KEYS], np.float64),
             h=h.download(), full=allv.download()[:sc.num_points], lo=ev.lo, hi=ev.hi)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_entry_world2_nccl(tmp_path, tiny_scene):
    from uavmvs_b200 import api, sharded
    if api.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    assert np.array_equal(r0["rec"].view(np.uint64), r1["rec"].view(np.uint64))        # every rank: the same bits
    assert np.all(r0["h"] == 3.0) and np.all(r1["h"] == 3.0)
    assert np.array_equal(r0["full"].view(np.uint32), r1["full"].view(np.uint32))
    sc = tiny_scene
    api.set_device(0)
    bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
    recs = []
    full = np.empty(sc.num_points, np.float32)
    for rank in range(2):
        lo, hi = sharded.shard_range(sc.num_points, rank, 2)
        e = _evaluate_unsharded(api, sc, bvh, sc.points[lo:hi], sc.normals[lo:hi])
        full[lo:hi] = e["recons"]
        recs.append([e["wsum64"], hi - lo, e["stats"][0], e["stats"][1], e["stats"][3], e["min"], e["max"], (hi - lo) * sc.num_views])
    want = sharded.fold_records(recs)
    got = dict(zip(api.SCORE_KEYS, r0["rec"]))
    assert int(got["points"]) == sc.num_points and int(got["rays"]) == want["stats"][0] and int(got["observations"]) == want["stats"][1]
    assert got["min"] == want["min"] and got["max"] == want["max"]
    assert abs(got["sum"] - want["sum"]) <= 1e-6 * abs(want["sum"])
    assert np.array_equal(r0["full"].view(np.uint32), full.view(np.uint32))
