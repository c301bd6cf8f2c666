"""Configs 3 and 4 of BASELINE.json at their full sizes on one B200, timed with CUDA events.

config 3  optimiser inner loop on the config-2 state (apps/optimize_trajectory/optimize_trajectory.cu:283-549):
          select mutually independent views (rule of :206-208, :290-310), remove them (K1 populate=false), K2, K3,
          7 candidate positions per selected view through batched K5 + K6 + device arg-min, re-add the moved
          views (K1), K2, K3, K8, K9.  Reported: latency per iteration and per stage, rays per iteration.
config 4  guidance volume (apps/generate_guidance_volume/generate_guidance_volume.cu:128-275): G^3 candidate
          positions on the reference's grid over the mesh AABB, those above the vertex height map kept
          (:139-192), batched K4 + K6; per voxel the 128x45 image's min/arg-min/max/arg-max are kept and a
          CRC32 of those four streams is printed (the images themselves would be 48 GB at 128^3).

Writes gpurun_out/<tag>_config3.json and gpurun_out/<tag>_config4.json.  A dev tool: nothing here is a bench line.
"""
import argparse, json, os, sys, time, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from uavmvs_b200 import api, scenes
import oracle  # icosphere + KD arrays only (host-side set-up, not timed, not the product path)

ap = argparse.ArgumentParser()
ap.add_argument("--tag", default="rXX")
ap.add_argument("--grid", type=int, default=128)
ap.add_argument("--batch", type=int, default=8192)
ap.add_argument("--limit", type=int, default=0, help="config 4: stop after this many positions (0 = all)")
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--full-eval", action="store_true", help="config 3: K2/K3/K8 over all points (the reference's launches) instead of the changed blocks")
ap.add_argument("--skip3", action="store_true")
ap.add_argument("--skip4", action="store_true")
args = ap.parse_args()
os.makedirs("gpurun_out", exist_ok=True)


def ev():
    return torch.cuda.Event(enable_timing=True)


sc = scenes.urban(1_000_000, 1_000_000, 2000, 256.0)
torch.cuda.init()
t0 = time.time()
bvh = api.BVHTree.build_sah(sc.verts, sc.faces)
build_s = time.time() - t0
rng = np.random.default_rng(4)
values = rng.uniform(0.0, 1.2, sc.num_points).astype(np.float32)
quals = rng.uniform(0.0, 1.0, sc.num_points).astype(np.float32)
sv = oracle.sphere_vertices(1.0, 3)
kn, kv = oracle.KD.build(sv).export()
kd = api.KDTree.create(kn, kv)
S = sv.shape[0]
N = sc.num_points
api.configure_heuristic()

# ------------------------------------------------------------------ config 3
if not args.skip3:
    cloud = api.PointCloud.create(sc.points, sc.normals)
    views = sc.views28()
    V = views.shape[0]
    obs = api.VectorArray.create(N, 32)
    recons, wrecons = api.Array.create(N), api.Array.create(N)

    def reeval():
        if args.full_eval:
            api.process_observation_rays(obs)
            api.evaluate_observation_rays(obs, recons)
            api.calculate_func_recons(recons, sc.target_recon, wrecons)
        else:
            api.reevaluate_observation_rays(obs, recons, sc.target_recon, wrecons)

    obs.null()
    api.update_observation_rays_batch(True, views, sc.max_distance, sc.width, sc.height, bvh, cloud, obs)
    reeval()
    score0 = api.reduce(wrecons)[0] / N
    gen = np.random.default_rng(7)
    iters_left = np.full(V, 100.0)

    def select(independence, want=16):
        """optimize_trajectory.cu:286-310: V draws from the remaining-iterations distribution, a draw is kept
        when no kept view is closer than sqrt(independence * 2) * max_distance."""
        min_sq = independence * 2.0 * sc.max_distance * sc.max_distance
        kept = []
        for idx in gen.choice(V, size=V, p=iters_left / iters_left.sum()):
            p = views[idx, :3]
            if all(float(((p - views[k, :3]) ** 2).sum()) >= min_sq for k in kept) and idx not in kept:
                kept.append(int(idx))
                if len(kept) == want:
                    break
        return kept

    independence = 1.0
    sel = select(independence)
    while len(sel) < 16 and independence > 0.05:   # the reference's --independence option "reduce[s the] constraint"
        independence *= 0.5
        sel = select(independence)
    hists = api.Array.create(len(sel) * 7 * S)
    rows = []
    for it in range(args.iters + 2):
        sel = select(independence)
        B = len(sel) * 7
        # regular-tetrahedron simplex around the view (:210-226) + 3 Nelder-Mead trial points = 7 candidates per view
        cand = (views[sel, None, :3] + gen.normal(0.0, 2.0, (len(sel), 7, 3))).reshape(-1, 3).astype(np.float32)
        e = [ev() for _ in range(6)]
        api.k1_counters()
        torch.cuda.synchronize()
        e[0].record()
        api.update_observation_rays_batch(False, views[sel], sc.max_distance, sc.width, sc.height, bvh, cloud, obs)
        e[1].record()
        reeval()
        e[2].record()
        hists.null()
        api.populate_spherical_histogram_batch(cand, sc.max_distance, bvh, cloud, kd, hists,
                                               target_recon=sc.target_recon, obs_rays=obs, recons=recons)
        e[3].record()
        vmin, imin, vmax, imax = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, kd, hists, B)
        e[4].record()
        # adopt the best candidate of every selected view (position only; orientation kept) and re-add
        best = vmin.reshape(len(sel), 7).argmin(1)
        moved = views[sel].copy()
        newpos = cand.reshape(len(sel), 7, 3)[np.arange(len(sel)), best]
        for k in range(len(sel)):
            R = moved[k, 3:19].reshape(4, 4)[:3, :3]
            moved[k, :3] = newpos[k]
            w2c = moved[k, 3:19].reshape(4, 4).copy()
            w2c[:3, 3] = -(R.astype(np.float64) @ newpos[k].astype(np.float64))
            moved[k, 3:19] = w2c.reshape(16)
        api.update_observation_rays_batch(True, moved, sc.max_distance, sc.width, sc.height, bvh, cloud, obs)
        reeval()
        score = api.reduce(wrecons)[0] / N
        e[5].record()
        torch.cuda.synchronize()
        views[sel] = moved
        iters_left[sel] -= 1
        rays = api.k1_counters()[3]
        ms = [e[i].elapsed_time(e[i + 1]) for i in range(5)]
        rows.append({"views": len(sel), "candidates": B, "rays": rays, "ms_total": e[0].elapsed_time(e[5]),
                     "ms_remove_K1": ms[0], "ms_K2_K3_K8": ms[1], "ms_K5_batch": ms[2], "ms_K6_argmin": ms[3],
                     "ms_readd_K1_K2_K3_K8_K9": ms[4], "score": score})
        print("config3 iter %d: %d views, %d candidates, %.2f ms (remove %.2f, K2+K3+K8 %.2f, K5 %.2f, K6 %.3f, re-add+eval %.2f), "
              "%d rays, score %.6f" % (it, len(sel), B, rows[-1]["ms_total"], *ms, rays, score), flush=True)
    timed = rows[2:]
    med = lambda k: float(np.median([r[k] for r in timed]))
    out3 = {"config": "3: optimiser inner loop on the config-2 state (1M triangles, 1M points, 2000 views resident)",
            "re_evaluation": "all points (reference launches)" if args.full_eval else "changed 32-point blocks only (uavmvs_reevaluate_observation_rays)",
            "independence": independence, "views_per_iteration": med("views"), "candidates_per_iteration": med("candidates"),
            "iterations_timed": len(timed), "ms_per_iteration_median": med("ms_total"),
            "stage_ms_median": {k: med(k) for k in timed[0] if k.startswith("ms_") and k != "ms_total"},
            "rays_per_iteration_median": med("rays"),
            "rays_per_s": med("rays") / med("ms_total") * 1e3, "score_before": score0, "score_after": rows[-1]["score"],
            "tree_build_s": build_s, "iterations": rows}
    json.dump(out3, open("gpurun_out/%s_config3%s.json" % (args.tag, "_full" if args.full_eval else ""), "w"), indent=1)
    print("config3:", json.dumps({k: v for k, v in out3.items() if k != "iterations"}), flush=True)
    del obs, recons, wrecons, hists, cloud

# ------------------------------------------------------------------ config 4
if not args.skip4:
    cloud = api.PointCloud.create(sc.points, sc.normals, values, quals)
    G = args.grid
    lo, hi = sc.verts.min(0), sc.verts.max(0)
    res = np.float32(max(hi[0] - lo[0], hi[1] - lo[1]) / G)      # one resolution for all three axes (:132-134)
    W = H = D = G
    # vertex height map (:139-151), ground level = lowest occupied cell (:153-167)
    hmap = np.full((H, W), -np.inf, np.float32)
    xi = np.minimum(((sc.verts[:, 0] - lo[0]) / res).astype(np.int64), W - 1)
    yi = np.minimum(((sc.verts[:, 1] - lo[1]) / res).astype(np.int64), H - 1)
    np.maximum.at(hmap, (yi, xi), sc.verts[:, 2])
    ground = float(hmap[np.isfinite(hmap)].min())
    hmap = np.where(np.isfinite(hmap), hmap - ground, 0.0).astype(np.float32)
    # sample positions (:174-192), y-major / x / z order; min_altitude 0
    ys, xs, zs = np.meshgrid(np.arange(H), np.arange(W), np.arange(D), indexing="ij")
    pz = (np.float32(ground) + zs.astype(np.float32) * res).astype(np.float32)
    keep = pz >= np.maximum(hmap[ys, xs], np.float32(0.0))
    px = ((xs.astype(np.float32) - res / np.float32(2.0)) * res + lo[0]).astype(np.float32)
    py = ((ys.astype(np.float32) - res / np.float32(2.0)) * res + lo[1]).astype(np.float32)
    pos = np.stack([px[keep], py[keep], pz[keep]], 1).astype(np.float32)
    if args.limit:
        sel = np.linspace(0, pos.shape[0] - 1, args.limit).astype(np.int64)
        pos = pos[sel]
    P = pos.shape[0]
    Bmax = args.batch
    hists = api.Array.create(Bmax * S)
    vmin_all, vmax_all = np.empty(P, np.float32), np.empty(P, np.float32)
    imin_all, imax_all = np.empty(P, np.uint32), np.empty(P, np.uint32)
    k4_ms = k6_ms = 0.0
    rays = 0
    api.k1_counters()
    torch.cuda.synchronize()
    w0 = time.time()
    for b0 in range(0, P, Bmax):
        p = pos[b0:b0 + Bmax]
        B = p.shape[0]
        e0, e1, e2 = ev(), ev(), ev()
        e0.record()
        hists.null()
        api.populate_spherical_histogram_batch(p, sc.max_distance, bvh, cloud, kd, hists)
        e1.record()
        r = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, kd, hists, B)
        e2.record()
        torch.cuda.synchronize()
        vmin_all[b0:b0 + B], imin_all[b0:b0 + B], vmax_all[b0:b0 + B], imax_all[b0:b0 + B] = r
        k4_ms += e0.elapsed_time(e1)
        k6_ms += e1.elapsed_time(e2)
        rays += api.k1_counters()[3]
        if (b0 // Bmax) % 16 == 0:
            print("config4: %d / %d positions, K4 %.1f s, K6 %.2f s, %d rays" % (b0 + B, P, k4_ms / 1e3, k6_ms / 1e3, rays), flush=True)
    wall = time.time() - w0
    crc = 0
    for a in (vmin_all, imin_all, vmax_all, imax_all):
        crc = zlib.crc32(a.tobytes(), crc)
    out4 = {"config": "4: guidance volume, %d^3 grid at %.3f m over the config-2 mesh AABB, 1M points, values U[0,1.2], qualities U[0,1] (seed 4)" % (G, float(res)),
            "grid": [W, H, D], "positions_in_free_space": int(keep.sum()), "positions_run": P, "batch": Bmax,
            "device_s_K4": k4_ms / 1e3, "device_s_K6_argminmax": k6_ms / 1e3, "wall_s": wall,
            "us_per_position": (k4_ms + k6_ms) * 1e3 / P, "rays": rays, "rays_per_s": rays / (k4_ms / 1e3),
            "pairs_per_s": P * float(N) / ((k4_ms + k6_ms) / 1e3),
            "positions_with_signal": int((vmax_all > 0).sum()), "max_of_max": float(vmax_all.max()),
            "argmax_of_max": int(vmax_all.argmax()), "crc32_min_argmin_max_argmax": "%08x" % crc,
            "tree_build_s": build_s}
    json.dump(out4, open("gpurun_out/%s_config4.json" % args.tag, "w"), indent=1)
    print("config4:", json.dumps(out4), flush=True)
