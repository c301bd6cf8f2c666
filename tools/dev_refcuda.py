"""Side reference on the GPU box: the reference's own kernels (oracle/_ref/libref_cuda_*.so, compiled from
/root/reference/libs/{eval,cacc} as they lie) against the oracle and the product on the same scenes, and the
reference's per-view launch loop timed on the same B200 (SURVEY.md section 8d "second side reference").
Developer / evidence tool; prints one JSON object per scene.

  python tools/dev_refcuda.py [--scenes tiny,dense,config1] [--config2 VIEWS]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import oracle
from uavmvs_b200 import api, scenes

ap = argparse.ArgumentParser()
ap.add_argument("--scenes", default="tiny,dense,config1")
ap.add_argument("--config2", type=int, default=0, help="time config 2 on this many of its 2000 views (0 = skip)")
ap.add_argument("--full", action="store_true", help="config 2: also compare rows/recons with the product (bit level)")
args = ap.parse_args()


def valid_equal(d_a, d_b, rows, max_rows=32):
    r = np.minimum(rows, max_rows)
    mask = np.arange(max_rows)[:, None] < r[None, :]
    a, b = d_a.view(np.uint32)[mask], d_b.view(np.uint32)[mask]
    return int((a != b).any(axis=-1).sum()) if a.ndim > 1 else int((a != b).sum())


def rel_err(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    den = np.maximum(np.maximum(np.abs(a), np.abs(b)), 1e-30)
    return float((np.abs(a - b) / den).max())


def compare(sc, name):
    out = {"scene": name, "points": sc.num_points, "views": sc.num_views, "tris": sc.num_tris}
    views = sc.views28()
    obvh = oracle.BVH.build(sc.verts, sc.faces, threads=0)
    ta = obvh.export()
    obs_o = oracle.ObsRays(sc.num_points, 32)
    oracle.update_observation_rays(obvh, True, views, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs_o)
    raw_rows_o, raw_o = obs_o.num_rows.copy(), obs_o.data.copy()
    oracle.process_observation_rays(obs_o)
    r_o = oracle.evaluate_observation_rays(obs_o)
    gbvh = api.BVHTree.create(ta.nodes, ta.aabbs, ta.tris, ta.indices)
    cloud = api.PointCloud.create(sc.points, sc.normals)
    obs = api.VectorArray.create(sc.num_points, 32)
    recons, wrecons = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
    api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height, gbvh, cloud, obs, recons, wrecons)
    d_g, rows_g = obs.download()
    r_g = recons.download()
    for kind in ("exact", "fast"):
        ref = oracle.RefCuda(sc.verts, sc.faces, sc.points, sc.normals, kind=kind)
        ref.update_observation_rays(True, views, sc.max_distance, sc.width, sc.height)
        rows_r, d_r = ref.download()
        o = {"observations": int(rows_r.sum()),
             "count_mismatch_vs_oracle": int((rows_r != raw_rows_o).sum()),
             "raw_row_mismatch_vs_oracle": valid_equal(raw_o[:, :sc.num_points], d_r, np.minimum(rows_r, raw_rows_o))}
        ref.process_observation_rays()
        rows_r, d_r = ref.download()
        o["sorted_row_mismatch_vs_oracle"] = valid_equal(obs_o.data[:, :sc.num_points], d_r, np.minimum(rows_r, obs_o.num_rows))
        o["sorted_row_mismatch_vs_product"] = valid_equal(d_g, d_r, np.minimum(rows_r, rows_g))
        o["count_mismatch_vs_product"] = int((rows_r != rows_g).sum())
        r_r = ref.evaluate_observation_rays()
        same = rows_r == obs_o.num_rows
        o["recon_rel_err_vs_oracle"] = rel_err(r_r[same], r_o[same])
        o["recon_rel_err_vs_product"] = rel_err(r_r[same], r_g[same])
        o["recon_flag_mismatch"] = int(((r_r == -1.0) != (r_o == -1.0)).sum())
        out[kind] = o
        del ref
    print(json.dumps(out), flush=True)


for name in [s for s in args.scenes.split(",") if s]:
    compare(getattr(scenes, name)(), name)

if args.config2:
    import torch
    sc = scenes.urban()
    V = min(args.config2, sc.num_views)
    sel = np.linspace(0, sc.num_views - 1, V).astype(np.int64)
    views = sc.views28()[sel]
    out = {"scene": sc.name, "views_timed": V, "points": sc.num_points, "tris": sc.num_tris}
    t = time.time()
    ref = oracle.RefCuda(sc.verts, sc.faces, sc.points, sc.normals, kind="fast")
    out["acc_build_and_upload_s"] = round(time.time() - t, 2)
    ms = [ref.time_evaluate(views, sc.max_distance, sc.width, sc.height, sc.target_recon) for _ in range(3)]
    rows_r, d_r = ref.download()
    out["reference_fast_ms"] = [round(x, 2) for x in ms]
    # the same views through the product (same tree topology: the reference's own SAH tree, exported)
    obvh = oracle.BVH.build(sc.verts, sc.faces, threads=0)
    ta = obvh.export()
    for label, bvh in (("product_on_acc_tree", api.BVHTree.create(ta.nodes, ta.aabbs, ta.tris, ta.indices)),
                       ("product_on_lbvh_tree", api.BVHTree.build_lbvh(sc.verts, sc.faces))):
        cloud = api.PointCloud.create(sc.points, sc.normals)
        obs = api.VectorArray.create(sc.num_points, 32)
        recons, wrecons = api.Array.create(sc.num_points), api.Array.create(sc.num_points)
        dviews = api.DeviceViews(views)
        stream = torch.cuda.current_stream()
        best = 1e9
        for it in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            s, mn, mx, st = api.evaluate_trajectory(dviews, sc.max_distance, sc.target_recon, sc.width, sc.height,
                                                    bvh, cloud, obs, recons, wrecons, stream=stream)
            e1.record(stream)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        out[label + "_ms"] = round(best, 3)
        out["rays"] = int(st[0])
        d_g, rows_g = obs.download()
        out[label + "_count_mismatch_vs_reference_fast"] = int((rows_g != rows_r).sum())
    out["reference_fast_grays_per_s"] = round(out["rays"] / min(ms) / 1e6, 4)
    out["product_grays_per_s"] = round(out["rays"] / out["product_on_lbvh_tree_ms"] / 1e6, 4)
    if args.full:
        refx = oracle.RefCuda(sc.verts, sc.faces, sc.points, sc.normals, kind="exact")
        out["reference_exact_ms"] = round(refx.time_evaluate(views, sc.max_distance, sc.width, sc.height, sc.target_recon), 2)
        rows_x, d_x = refx.download()
        out["product_count_mismatch_vs_reference_exact"] = int((rows_g != rows_x).sum())
        out["product_sorted_row_mismatch_vs_reference_exact"] = valid_equal(d_g, d_x, np.minimum(rows_x, rows_g))
        r_x = refx.evaluate_observation_rays()
        same = rows_x == rows_g
        out["product_recon_rel_err_vs_reference_exact"] = rel_err(r_x[same], recons.download()[same])
    print(json.dumps(out), flush=True)
