#!/usr/bin/env python
"""bench.py -- headline benchmark of the visibility-and-reconstructability evaluation.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config {2,3,4,5}]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

--config 2 (default, BASELINE.json configs[1]): synthetic 1M-triangle urban block, 1M proxy points per GPU, 2k-view
           trajectory; a "step" is one full evaluation (clear obs, K1 over all views, K2, K3+K8, K9, score collective)
           through uavmvs_evaluate_trajectory_sharded.  Weak scaling: rank r evaluates its own 1M-point sample (seed
           + r) of the SAME block against the replicated mesh, so the per-GPU workload is the same at every N; the
           trajectory score is combined by one ncclAllGather of a 64-byte record per rank INSIDE the timed call.
           On one GPU the line also carries `extra` keys measured in the same run (configs 3 and 4 on bounded
           samples); at every N (unless --no-strong) a `strong` block: config 5 at its fixed total size, sharded
           over the N ranks, so the driver's 1/2/4/8 runs hold the strong-scaling curve as well.
--config 3 optimiser inner loop on the resident config-2 state (step = one iteration).
--config 4 guidance volume, 128^3 grid (step = one batch of 8192 candidate positions; --positions bounds the run).
--config 5 10M-triangle city, 8M points in total sharded over the ranks, 10k views (strong scaling).

  value  occlusion rays/s, whole job, inputs resident in HBM, CUDA events on the launching stream, max over ranks
  e2e    the same through the C-ABI call with HOST buffers: trajectory H2D from pinned memory, evaluation,
         collective, D2H of score + per-point recon + per-point counts into pinned memory
  --impl reference   the CPU path (oracle port of libs/eval math over the reference's acc tree walk) on the host
         cores with an explicit OpenMP team, a bounded sample of the same workload per step

One JSON line on stdout (rank 0).  Nothing here reads /root/reference.  Non-zero exit when the in-run parity check
against the oracle fails.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "occlusion_rays_per_s"
UNIT = "rays/s"
CHUNK = 1024          # default --shard-chunk: points per interleaved chunk of the strong-scaling split (= one 1024-point super-block)
L2_NOTE = "GPU arm: L2 flushed between steps (256 MiB fill, outside the events); CPU arm: n/a"
CONFIGS = {
    2: dict(tris=1_000_000, points=1_000_000, views=2000, extent=256.0, scaling="weak"),
    3: dict(tris=1_000_000, points=1_000_000, views=2000, extent=256.0, scaling="weak"),
    4: dict(tris=1_000_000, points=1_000_000, views=2000, extent=256.0, scaling="weak"),
    5: dict(tris=10_000_000, points=8_000_000, views=10_000, extent=800.0, scaling="strong"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5])
    ap.add_argument("--tris", type=int, default=0)
    ap.add_argument("--points", type=int, default=0, help="proxy points per GPU (config 2-4) / in total (config 5)")
    ap.add_argument("--views", type=int, default=0)
    ap.add_argument("--extent", type=float, default=0.0)
    ap.add_argument("--tree", default="sah", choices=["sah", "ploc", "lbvh"], help="GPU tree build (product path)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU work budget of the cpu_baseline sample")
    ap.add_argument("--no-extras", action="store_true", help="config 2: skip the config-3 / config-4 extra keys")
    ap.add_argument("--no-strong", action="store_true", help="config 2: skip the config-5 strong-scaling block")
    ap.add_argument("--strong-steps", type=int, default=5)
    ap.add_argument("--positions", type=int, default=0, help="config 4: number of candidate positions (0 = steps x 8192)")
    ap.add_argument("--grid", type=int, default=128, help="config 4: voxels per axis")
    ap.add_argument("--shard-chunk", type=int, default=CHUNK, help="strong scaling: C = interleaved C-point Z-order chunks (default 1024), 0 = contiguous ranges per rank")
    a = ap.parse_args()
    c = CONFIGS[a.config]
    a.tris, a.points, a.views = a.tris or c["tris"], a.points or c["points"], a.views or c["views"]
    a.extent, a.scaling = a.extent or c["extent"], c["scaling"]
    return a


def config_dict(a, world):
    """The same dict in both arms (the driver compares them)."""
    what = {2: "full evaluation K1+K2+K3+K8+K9 + score collective",
            3: "optimiser iteration (remove 16 views, K2+K3+K8, one Nelder-Mead step per view on the device = 96 objective "
               "evaluations K5+K6+arg-min, re-add, K2+K3+K8+K9)",
            4: "guidance volume %d^3 grid, batched K4+K6+arg-min/max" % a.grid,
            5: "full evaluation K1+K2+K3+K8+K9 + score collective, points sharded over the GPUs"}[a.config]
    per = "%d proxy points per GPU x %d GPU" % (a.points, world) if a.scaling == "weak" else \
        "%d proxy points in total over %d GPU" % (a.points, world)
    return {"workload": "config%d: synthetic urban block %.0f m, %d-triangle mesh, %s, %d views, max_distance 80, "
                        "1920x1080, %s" % (a.config, a.extent, a.tris, per, a.views, what),
            "config": a.config, "tris": a.tris, "points": a.points, "views": a.views, "extent": a.extent,
            "n_gpus": world, "l2": L2_NOTE}


# --------------------------------------------------------------------------- scenes and shards
def make_scene(a):
    from uavmvs_b200 import scenes
    n = a.points if a.scaling == "strong" else 1000          # weak: the per-rank clouds are sampled separately
    return scenes.urban(a.tris, n, a.views, a.extent, seed=1)


def rank_cloud(a, sc, rank, world):
    """This rank's proxy points.  weak: an independent area-weighted sample of the SAME mesh per rank (rank 0's is
    the N = 1 cloud), so every rank has the N = 1 workload.  strong: interleaved --shard-chunk C-point chunks of the Z-ordered
    cloud, every run of `world` consecutive chunks giving one chunk to each rank with a rotating owner; C = 0: contiguous
    ranges of ceil(N / G) points (SURVEY.md 8e's partition).  Measured on config 5 at N = 8 (profiles/r02_fin3_strong_split.txt):
    1024-point chunks 3.198 ms, 8192-point chunks 3.206 ms, contiguous ranges 3.208 ms (each rank then works on one region of
    the city and of the tree, k_trace 1.38 - 1.62 ms, but the ranks hold 11.05 M to 12.79 M rays and the slowest sets the
    time).  The split is not what limits strong scaling; the default stays 1024."""
    from uavmvs_b200 import scenes
    if a.scaling == "weak":
        return scenes.sample_cloud(sc.verts, sc.faces, a.points, 2 + 1000 * rank)
    n = sc.num_points
    if world == 1:
        return sc.points, sc.normals
    chunk = int(getattr(a, "shard_chunk", CHUNK))
    if chunk <= 0:
        per = ((n + world - 1) // world + 31) // 32 * 32
        lo, hi = min(n, rank * per), min(n, (rank + 1) * per)
        return np.ascontiguousarray(sc.points[lo:hi]), np.ascontiguousarray(sc.normals[lo:hi])
    chunks = np.arange((n + chunk - 1) // chunk, dtype=np.uint64)
    rot = ((chunks // np.uint64(world)) * np.uint64(2654435761) >> np.uint64(7)) % np.uint64(world)
    mine = chunks[(chunks + rot) % np.uint64(world) == np.uint64(rank)].astype(np.int64)
    idx = (mine[:, None] * chunk + np.arange(chunk)[None, :]).reshape(-1)
    idx = idx[idx < n]
    return np.ascontiguousarray(sc.points[idx]), np.ascontiguousarray(sc.normals[idx])


# --------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._stop_evt.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "nvml unavailable"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def visible_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:  # noqa: BLE001
            return local_rank
    return local_rank



# --------------------------------------------------------------------------- CPU arm (oracle/)
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        return os.cpu_count() or 1


def cpu_team(orc):
    """Sets the oracle's OpenMP team to every core this process may use and returns the team size a parallel region
    really gets (torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the affinity mask says nothing
    about that)."""
    return orc.set_threads(host_threads())


def cpu_tree(orc, sc, threads):
    """The reference's CPU tree: acc::BVHTree built by oracle/_ref (the reference's own code)
    when that library travelled with the repo, else by the bit-identical restatement."""
    t0 = time.perf_counter()
    if orc.ref() is not None:
        arrays = orc.RefBVH(sc.verts, sc.faces, threads=threads).export()
        tree, built_by = orc.BVH.from_arrays(arrays), "oracle/_ref acc::BVHTree (reference code)"
    else:
        tree, built_by = orc.BVH.build(sc.verts, sc.faces, threads=threads), "oracle port of acc::BVHTree"
    return tree, built_by, time.perf_counter() - t0


def cpu_step(orc, tree, sc, views, pts, nrm):
    """One CPU evaluation over a set of views: K1 (OpenMP over points, views in order), K2, K3,
    K8, K9 -- the reference's algorithm, per-pair cull included."""
    obs = orc.ObsRays(pts.shape[0], 32)
    t0 = time.perf_counter()
    _, cnt = orc.update_observation_rays(tree, True, views, sc.max_distance, sc.width, sc.height, pts, nrm, obs,
                                         want_counters=True)
    t1 = time.perf_counter()
    orc.process_observation_rays(obs)
    recons = orc.evaluate_observation_rays(obs)
    w = orc.calculate_func_recons(recons, sc.target_recon)
    red = orc.reduce(w)
    t2 = time.perf_counter()
    return {"rays": int(cnt[0]), "pairs": int(cnt[5]), "counters": [int(x) for x in cnt], "seconds": t2 - t0,
            "k1_seconds": t1 - t0, "score": float(red[0]) / pts.shape[0], "obs": obs, "recons": recons}


def pick_views(num_views, count):
    count = max(1, min(num_views, count))
    return np.unique(np.linspace(0, num_views - 1, count).round().astype(int))


def calibrate_cpu_sample(orc, tree, sc, pts, nrm, budget_s):
    """Number of (evenly strided) views whose CPU evaluation takes about budget_s seconds."""
    probe = pick_views(sc.num_views, 16)
    r = cpu_step(orc, tree, sc, sc.views28()[probe], pts, nrm)
    per_view = max(r["seconds"] / len(probe), 1e-6)
    return int(max(16, min(sc.num_views, budget_s / per_view)))


def bytes_per_ray(counters):
    """B_ray = 80 n_inner + 16 n_leaf + 48 n_tri + 4 n_hit, counters of the oracle's
    instrumented reference traversal on the same tree."""
    rays, inner, leaf, tri, hit = counters[:5]
    if rays == 0:
        return 0.0, {}
    per = {"n_inner": inner / rays, "n_leaf": leaf / rays, "n_tri": tri / rays, "n_hit": hit / rays}
    return 80.0 * per["n_inner"] + 16.0 * per["n_leaf"] + 48.0 * per["n_tri"] + 4.0 * per["n_hit"], per


def reference_cloud(a, sc, world):
    """The points of the whole job (all ranks' shares), as the GPU arm evaluates them at this N."""
    if a.scaling == "strong" or world == 1:
        return rank_cloud(a, sc, 0, 1)
    parts = [rank_cloud(a, sc, r, world) for r in range(world)]
    return np.concatenate([p for p, _ in parts]), np.concatenate([n for _, n in parts])


def run_reference(a, rank, world):
    if rank != 0:
        return
    import oracle as orc
    orc.build()
    team = cpu_team(orc)
    sc = make_scene(a)
    tree, built_by, build_s = cpu_tree(orc, sc, team)
    total = max(1, a.steps + a.warmup)
    budget = min(12.0, 150.0 / total)                 # whole run within a few minutes
    if a.config == 4:
        return run_reference_config4(a, orc, sc, tree, built_by, build_s, team, budget, world)
    pts, nrm = reference_cloud(a, sc, world)
    nviews = calibrate_cpu_sample(orc, tree, sc, pts, nrm, budget)
    vidx = pick_views(sc.num_views, nviews)
    views = sc.views28()[vidx]
    for _ in range(a.warmup):
        cpu_step(orc, tree, sc, views, pts, nrm)
    secs, rays, pairs = 0.0, 0, 0
    for _ in range(a.steps):
        r = cpu_step(orc, tree, sc, views, pts, nrm)
        secs += r["seconds"]; rays += r["rays"]; pairs += r["pairs"]
    value = rays / secs
    sample = ("%d of %d views (evenly strided) x all %d points of the job per step; tree: %s (built in %.2f s, not "
              "timed); kernel math: oracle port of libs/eval/kernels.cu (the reference has no CPU implementation of "
              "it), OpenMP over points, team of %d threads" % (len(vidx), sc.num_views, pts.shape[0], built_by, build_s, team))
    if a.config == 3:
        sample += "; config 3 is represented on the CPU by the evaluation it repeats (K1 + K2 + K3 + K8 + K9)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * secs / a.steps, "higher_is_better": True, "scaling": a.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(a, world),
            "same_workload": a.config != 3, "pairs_per_s": pairs / secs,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": team, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def run_reference_config4(a, orc, sc, tree, built_by, build_s, team, budget, world):
    """Guidance volume on the CPU: per candidate position K4 over all points + K6 (oracle port), OpenMP over points."""
    from uavmvs_b200 import apps
    rng = np.random.default_rng(4)
    pts, nrm = rank_cloud(a, sc, 0, 1)
    values = rng.uniform(0.0, 1.2, pts.shape[0]).astype(np.float32)
    quals = rng.uniform(0.0, 1.0, pts.shape[0]).astype(np.float32)
    sv = orc.sphere_vertices(1.0, 3)
    kd = orc.KD.build(sv)
    pos, free_total, res = apps.GuidanceVolume._positions(sc.verts, a.grid)

    def one(p):
        h, bins, _ = orc.populate_spherical_histogram(tree, kd, p, sc.max_distance, pts, nrm, values, quals, sv.shape[0])
        orc.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, sv, h.astype(np.float32))
        return int((bins != 0xFFFFFFFF).sum())            # visible rays (a lower bound of the rays cast)
    t0 = time.perf_counter()
    probe = pos[np.linspace(0, pos.shape[0] - 1, 8).astype(np.int64)]
    for p in probe:
        one(p)
    per = max((time.perf_counter() - t0) / len(probe), 1e-6)
    npos = int(max(4, min(pos.shape[0], budget / per)))
    sel = pos[np.linspace(0, pos.shape[0] - 1, npos).astype(np.int64)]
    secs, rays = 0.0, 0
    for it in range(a.warmup + a.steps):
        t0 = time.perf_counter()
        r = 0
        for p in sel:
            r += one(p)
        if it >= a.warmup:
            secs += time.perf_counter() - t0; rays += r
    value = rays / secs if secs > 0 else 0.0
    sample = ("%d of %d free-space positions (evenly strided) x all %d points per step; tree: %s (%.2f s, not timed); "
              "oracle port of K4 + K6, OpenMP over points, team of %d threads; rays = visible rays (the oracle's K4 "
              "does not count occluded ones)" % (npos, free_total, pts.shape[0], built_by, build_s, team))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * secs / a.steps, "higher_is_better": True, "scaling": a.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(a, world),
            "us_per_position": 1e6 * secs / (a.steps * npos),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": team, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# --------------------------------------------------------------------------- B200 arm
def static_ncu():
    """Per-launch counters of the dominant kernel from the committed `ncu --set full` capture (profiles/): they cannot
    be measured outside a profiler, so they are STATIC and labelled with the kernel version they belong to."""
    for name in ("k1_static_r02.json", "k1_traffic.json"):
        p = os.path.join(ROOT, "profiles", name)
        try:
            d = json.load(open(p))
            d["file"] = "profiles/" + name
            return d
        except Exception:  # noqa: BLE001
            continue
    return None


class Timer:
    """CUDA events on the launching stream; L2 flushed before every step, outside the events."""

    def __init__(self, torch, stream, flush):
        self.torch, self.stream, self.flush = torch, stream, flush

    def run(self, step_fn, steps, after=None):
        ms, last = 0.0, None
        for _ in range(steps):
            self.flush.fill_(1)
            e0, e1 = self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True)
            e0.record(self.stream)
            last = step_fn()
            e1.record(self.stream)
            e1.synchronize()
            ms += e0.elapsed_time(e1)
            if after:
                after()
        return ms, last


def parity_check(orc, api, sc, bvh, cloud, pts, nrm, nviews=24):
    """The oracle walks the exported tree of the timed run over `nviews` evenly strided views x this rank's points;
    the GPU evaluates the same views: rays, counters and stored rows must be bit-identical, recon within 1e-5."""
    vidx = pick_views(sc.num_views, nviews)
    views = np.ascontiguousarray(sc.views28()[vidx])
    nodes, aabbs, tris, indices = bvh.export()
    tree = orc.BVH.from_arrays(orc.TreeArrays(nodes, aabbs, tris, indices))
    probe = cpu_step(orc, tree, sc, views, pts, nrm)
    n = pts.shape[0]
    obs = api.VectorArray.create(n, 32)
    rc, wr = api.Array.create(n), api.Array.create(n)
    s_sum, _, _, stats = api.evaluate_trajectory(views, sc.max_distance, sc.target_recon, sc.width, sc.height, bvh, cloud, obs, rc, wr)
    data, rows = obs.download()
    r_g = rc.download()
    o = probe["obs"]
    mask = np.arange(32)[:, None] < np.minimum(rows, 32)[None, :]
    counts_equal = bool(np.array_equal(rows, o.num_rows))
    rows_equal = counts_equal and bool(np.array_equal(o.data.view(np.uint32)[mask], data.view(np.uint32)[mask]))
    r_o = probe["recons"]
    diff = np.abs(r_g.astype(np.float64) - r_o)
    mag = np.maximum(np.abs(r_g), np.abs(r_o)).astype(np.float64)
    big = mag > 1e-3                                      # recon of a well-conditioned point (see the gate below)
    gate = 1e-5 * mag + 1e-6                              # SURVEY.md 7E, per element
    res = {"views": int(len(vidx)), "points": int(n), "rays_gpu": int(stats[0]), "rays_oracle": int(probe["rays"]),
           "rays_equal": int(stats[0]) == int(probe["rays"]), "counts_equal": counts_equal, "rows_equal": rows_equal,
           "recon_gate": "|a - b| <= 1e-5 max(|a|, |b|) + 1e-6 per point (SURVEY.md 7E: 1 - logistic cancels for the wide-parallax "
                         "pairs of a point seen by two or three views, so recon ~ 1e-6 carries an absolute error)",
           "recon_gate_violations": int((diff > gate).sum()),
           "recon_max_rel_above_1e-3": float((diff[big] / mag[big]).max()) if big.any() else 0.0,
           "recon_max_abs": float(diff.max()),
           "tree": "the timed run's own tree, exported through uavmvs_bvh_export"}
    res["ok"] = bool(res["rays_equal"] and counts_equal and rows_equal and res["recon_gate_violations"] == 0)
    probe.pop("obs"); probe.pop("recons")
    return res, probe, (nodes.nbytes + aabbs.nbytes + tris.nbytes)


def gather_rank_stats(torch, dist, world, vec):
    mine = torch.tensor(vec, dtype=torch.float64, device="cuda")
    if world == 1:
        return [mine.tolist()]
    every = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    return [e.tolist() for e in every]


def evaluation_block(a, torch, dist, api, comm, rank, world, stream, flush, steps, warmup, want_e2e=True):
    """Times the sharded evaluation of configs 2 / 5 on this rank's shard.  Returns a dict of measurements plus the
    handles (for the parity check)."""
    sc = make_scene(a)
    pts, nrm = rank_cloud(a, sc, rank, world)
    n_local = pts.shape[0]
    views_np = sc.views28()
    V = views_np.shape[0]
    t0 = time.perf_counter()
    bvh = getattr(api.BVHTree, "build_" + a.tree)(sc.verts, sc.faces)
    api.synchronize()
    build_s = time.perf_counter() - t0
    cloud = api.PointCloud.create(pts, nrm)
    obs = api.VectorArray.create(n_local, 32)
    recons, wrecons = api.Array.create(n_local), api.Array.create(n_local)
    dviews = api.DeviceViews(views_np)
    h_views = torch.from_numpy(views_np.copy()).pin_memory()
    h_recons = torch.empty(n_local, dtype=torch.float32).pin_memory()
    h_rows = torch.empty(n_local, dtype=torch.int32).pin_memory()
    h_recons_np, h_rows_np = h_recons.numpy(), h_rows.numpy().view(np.uint32)
    args = (sc.max_distance, sc.target_recon, sc.width, sc.height, bvh, cloud, obs, recons, wrecons)

    def resident_step():
        return api.evaluate_trajectory_sharded(comm, dviews, *args, stream=stream)

    def e2e_step():
        # host trajectory -> C ABI (H2D inside) -> evaluation -> collective -> D2H of score, per-point recon and counts
        return api.evaluate_trajectory_sharded(comm, h_views.numpy(), *args, h_recons=h_recons_np, h_num_rows=h_rows_np,
                                               stream=stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    timer = Timer(torch, stream, flush)
    api.profile_k1_enable(True)
    for _ in range(warmup):
        flush.fill_(1)
        resident_step()
    k_ms, k1_ms, coll_ms = np.zeros(4), np.zeros(5), [0.0]

    def after():
        k_ms[:] += np.array(api.profile_last())
        k1_ms[:] += np.array(api.profile_k1())
        coll_ms[0] += api.profile_last_collective()

    sampler = ClockSampler(visible_gpu_index(int(os.environ.get("LOCAL_RANK", "0"))))
    barrier()
    api.launch_count(reset=True)
    api.k1_counters(reset=True)
    sampler.start()
    ms, last = timer.run(resident_step, steps, after)
    barrier()
    clocks = sampler.stop()
    launches = api.launch_count(reset=True)
    score, _ = last
    e2e_ms, e2e_ok = None, None
    if want_e2e:
        for _ in range(min(warmup, 3)):
            e2e_step()
        barrier()
        e2e_ms, _ = timer.run(e2e_step, steps)
        # the host buffers the timed e2e steps filled hold the device results (checked outside the timed region)
        e2e_ok = bool(np.array_equal(h_recons_np.view(np.uint32), recons.download().view(np.uint32))
                      and np.array_equal(h_rows_np, obs.download(data=False)[1]))
        barrier()
    api.profile_k1_enable(False)
    bits = {"hw_slowdown": 1, "hw_thermal_slowdown": 2, "sw_thermal_slowdown": 4, "sw_power_cap": 8}
    rays_local = api.k1_counters(reset=True)[3] / max(steps, 1)
    per_rank = gather_rank_stats(torch, dist, world, [ms / steps, float(k_ms[0]) / steps, float(k1_ms[3]) / steps,
                                                      float(n_local), clocks["sm_mhz"] or 0.0,
                                                      float(sum(bits[r] for r in clocks["reasons"])),
                                                      float(k1_ms[0]) / steps, float(k1_ms[1]) / steps, float(k1_ms[2]) / steps,
                                                      float(k1_ms[4]) / steps, float(k_ms[2]) / steps, float(rays_local)])
    if world > 1:
        t = torch.tensor([ms, e2e_ms or 0.0], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_max = t.tolist()
        e2e_ms = e2e_max if want_e2e else None
        mask = 0
        for r in per_rank:
            mask |= int(r[5])
        mhz = [r[4] for r in per_rank]
        clocks = dict(clocks, sm_mhz=min(mhz), sm_mhz_per_rank=mhz, reasons=sorted(r for r, b in bits.items() if mask & b))
    return dict(sc=sc, pts=pts, nrm=nrm, bvh=bvh, cloud=cloud, n_local=n_local, V=V, build_s=build_s, ms=ms, steps=steps,
                e2e_ms=e2e_ms, e2e_ok=e2e_ok, score=score, k_ms=k_ms / steps, k1_ms=k1_ms / steps, coll_ms=coll_ms[0] / steps,
                clocks=clocks, launches=launches, per_rank=[r[:5] + r[6:] for r in per_rank], views_bytes=int(views_np.nbytes))


def roofline(blk, probe, tree_bytes, hbm_peak, peak_src, sms=148):
    """The roof that binds the dominant kernel (k_trace) is instruction issue: achieved = warp instructions of one launch
    (static, from the committed ncu capture of the same workload) / the kernel's live duration; peak = one warp
    instruction per cycle and SM sub-partition at the SM clock sampled during the run.  The HBM figure the contract asks
    for (algorithmic bytes per SURVEY.md 8(d) / HBM peak) is kept beside it and labelled: it is not a bound here."""
    st = static_ncu() or {}
    kt = st.get("k_trace_ncu", {})
    score, V, n_local = blk["score"], blk["V"], blk["n_local"]
    world = max(1, len(blk["per_rank"]))
    rays, observations = score["rays"], score["observations"]
    b_ray, per_ray = bytes_per_ray(probe["counters"])
    stored = min(observations, 32 * score["points"])
    k1_ms, trace_ms = float(blk["k_ms"][0]), float(blk["k1_ms"][3])
    alg_bytes = (rays / world) * b_ray + n_local * (32 + 4) + (stored / world) * 16 + V * 112
    hbm_achieved = alg_bytes / (k1_ms / 1e3) / 1e9 if k1_ms > 0 else 0.0
    mhz = blk["clocks"].get("sm_mhz") or blk["clocks"].get("sm_max_mhz") or 1965.0
    inst = kt.get("warp_instructions")
    peak_ginst = sms * 4 * mhz * 1e6 / 1e9
    achieved_ginst = (inst / (trace_ms / 1e3) / 1e9) if (inst and trace_ms > 0) else None
    dram = st.get("dram_bytes_per_launch")
    return {"bound": "issue", "kernel": "k_trace (persistent 4-wide BVH traversal, %.0f %% of the step)" %
            (100.0 * trace_ms / max(blk["ms"] / blk["steps"], 1e-9)),
            "achieved": achieved_ginst, "peak": peak_ginst, "unit": "Gwarp-inst/s",
            "frac": None if achieved_ginst is None else achieved_ginst / peak_ginst,
            "kernel_ms": trace_ms, "k1_ms": k1_ms,
            "k1_kernels_ms": dict(zip(("k_view_masks", "caps_scan", "k_cull_emit_q", "k_trace", "k_retire"),
                                      (float(x) for x in blk["k1_ms"]))),
            "peak_source": "148 SMs x 4 schedulers x 1 warp instruction / cycle x %.0f MHz (median SM clock sampled during the timed region)" % mhz,
            "traffic": dram,
            "static": {"what": "warp_instructions, traffic (DRAM bytes of the K1 kernels per launch), ncu_k_trace: one `ncu --set full` "
                               "capture of this workload, NOT measured in this run", "file": st.get("file"),
                       "kernel_version": st.get("kernel_version"), "ncu_k_trace": kt or None},
            "dram": None if not dram else {"bytes_per_launch": dram, "gbs": dram / (k1_ms / 1e3) / 1e9 if k1_ms > 0 else None,
                                            "frac_of_hbm_peak": dram / (k1_ms / 1e3) / 1e9 / hbm_peak if k1_ms > 0 else None},
            "l2": st.get("l2"),
            "hbm_algorithmic": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                                "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes, "bytes_per_ray": b_ray,
                                "per_ray": per_ray,
                                "note": "SURVEY.md 8(d) bytes of the REFERENCE's access pattern (oracle's instrumented walk of the same "
                                        "exported tree, parity-check sample) / K1 time / HBM peak.  Not a bound: the compressed tree "
                                        "(%.0f MB in the reference layout) is L2-resident, DRAM traffic is the ray records and "
                                        "observation rows only, so this ratio can exceed 1" % (tree_bytes / 1e6)}}


def extras_config34(a, torch, api, stream, blk, line):
    """Configs 3 and 4 on the resident config-2 scene, bounded samples, same process (driver-visible record)."""
    from uavmvs_b200 import apps
    sc, bvh = blk["sc"], blk["bvh"]
    pts, nrm = blk["pts"], blk["nrm"]
    scn = type(sc)(**{**sc.__dict__, "points": pts, "normals": nrm, "values": np.zeros(len(pts), np.float32),
                      "qualities": np.ones(len(pts), np.float32)})
    sv, kd = apps.sphere_kdtree()
    S = sv.shape[0]
    ev = lambda: torch.cuda.Event(enable_timing=True)     # noqa: E731
    out = {}
    # ---- config 3: optimiser iterations
    loop = apps.OptimizerLoop(scn, bvh, blk["cloud"], kd, S)
    loop.populate(stream)
    loop.relax_independence()
    times, stage = [], []
    order = ["remove", "reeval", "k5", "k6", "readd", "end"]
    for it in range(8):
        marks = {}

        def mark(name, marks=marks):
            e = ev(); e.record(stream); marks[name] = e
        api.k1_counters()
        r = loop.iteration(stream, mark)
        torch.cuda.synchronize()
        ms = [marks[order[i]].elapsed_time(marks[order[i + 1]]) for i in range(5)]
        if it >= 2:
            times.append(marks["remove"].elapsed_time(marks["end"]))
            stage.append(ms + [api.k1_counters()[3], r["views"], r["candidates"]])
    st = np.median(np.array(stage), axis=0)
    out["config3_ms_per_iteration"] = float(np.median(times))
    out["config3"] = {"iterations_timed": len(times), "views_per_iteration": int(st[6]), "candidates_per_iteration": int(st[7]),
                      "rays_per_iteration": int(st[5]), "independence": loop.independence,
                      "stage_ms": dict(zip(("remove_K1", "K2_K3_K8_changed_blocks", "K5_batch", "K6_argmin", "readd_K1_K2_K3_K8_K9"),
                                           (float(x) for x in st[:5]))),
                      "what": "optimize_trajectory.cu:283-549 on the resident config-2 state, median of %d iterations; candidate "
                              "positions are random trial points around the removed views" % len(times)}
    del loop
    # ---- config 3 with the simplex-downhill step itself on the device (uavmvs_nm_step): the full optimiser iteration
    loop = apps.OptimizerLoop(scn, bvh, blk["cloud"], kd, S, nelder_mead=True, min_distance=2.5)
    loop.populate(stream)
    loop.relax_independence()
    times, stage = [], []
    order = ["remove", "reeval", "nm", "readd", "end"]
    for it in range(8):
        marks = {}

        def mark(name, marks=marks):
            e = ev(); e.record(stream); marks[name] = e
        api.k1_counters()
        r = loop.iteration_nm(stream, mark)
        torch.cuda.synchronize()
        ms = [marks[order[i]].elapsed_time(marks[order[i + 1]]) for i in range(4)]
        if it >= 2:
            times.append(marks["remove"].elapsed_time(marks["end"]))
            stage.append(ms + [api.k1_counters()[3], r["views"]])
    st = np.median(np.array(stage), axis=0)
    out["config3_nm_ms_per_iteration"] = float(np.median(times))
    out["config3_nm"] = {"iterations_timed": len(times), "views_per_iteration": int(st[5]), "rays_per_iteration": int(st[4]),
                         "objective_evaluations_per_view": 6, "score": r["score"],
                         "stage_ms": dict(zip(("remove_K1", "K2_K3_K8_changed_blocks", "nelder_mead_step_on_device", "readd_K1_K2_K3_K8_K9_and_D2H"),
                                              (float(x) for x in st[:4]))),
                         "what": "the same iteration with libs/opti/nelder_mead.h's step for the 16 selected views on the device "
                                 "(uavmvs_nm_step: 4 + 1 + 1 objective evaluations per view = altitude / closest_point airspace test, K5, "
                                 "K6, arg-min, in three batches, decisions by device kernels); H2D: the selected views; no D2H until "
                                 "the score and the 16 moved poses at the end of the iteration"}
    del loop
    # ---- config 4: a fixed 64 k-position sample of the 128^3 volume
    rng = np.random.default_rng(4)
    cloud4 = api.PointCloud.create(pts, nrm, rng.uniform(0.0, 1.2, len(pts)).astype(np.float32),
                                   rng.uniform(0.0, 1.0, len(pts)).astype(np.float32))
    vol = apps.GuidanceVolume(scn, bvh, cloud4, kd, S, grid=a.grid, batch=8192)
    npos = vol.subsample(65536)
    vol_w = apps.GuidanceVolume(scn, bvh, cloud4, kd, S, grid=a.grid, batch=8192)
    vol_w.subsample(8192)
    vol_w.run(stream)                                   # warm-up (bit matrix, first buffers)
    # two identical runs under the clock sampler: the first lets the ray buffers grow to the largest batch of THIS sample (the
    # library reallocates them inside a call when a batch needs more), the second is the reported one.  This block keeps
    # k_trace busy for ~1 s, unlike the 4 ms steps above: sustained clocks and throttle reasons are part of the record
    torch.cuda.synchronize()
    sampler4 = ClockSampler(visible_gpu_index(int(os.environ.get("LOCAL_RANK", "0"))))
    sampler4.start()
    f0, f1 = ev(), ev()
    f0.record(stream)
    r1 = vol.run(stream)
    f1.record(stream)
    torch.cuda.synchronize()
    api.k1_counters()
    e0, e1 = ev(), ev()
    e0.record(stream)
    r2 = vol.run(stream)
    e1.record(stream)
    torch.cuda.synchronize()
    clocks4 = sampler4.stop()
    rays4 = api.k1_counters()[3]
    ms4 = e0.elapsed_time(e1)
    out["config4_us_per_position"] = 1e3 * ms4 / npos
    out["config4"] = {"positions": npos, "of_free_space_positions": vol.free_total, "grid": a.grid, "ms": ms4, "rays": int(rays4),
                      "rays_per_s": rays4 / (ms4 / 1e3), "crc32_min_argmin_max_argmax": r1["crc32"],
                      "reproducible": r1["crc32"] == r2["crc32"], "positions_with_signal": int((r1["max"] > 0).sum()),
                      "ms_first_run": f0.elapsed_time(f1), "clocks": clocks4,
                      "what": "generate_guidance_volume.cu:128-275, evenly strided sample of the free-space voxels of the 128^3 grid "
                              "x all points (values U[0,1.2], qualities U[0,1], seed 4), batches of 8192, wall of the event pair "
                              "(includes the per-batch result copies); second of two identical runs, so that no buffer grows "
                              "inside the timed region"}
    del vol, vol_w, cloud4
    # ---- K7 estimate_capture_difficulty on a 2^18-point sample of the cloud: wide-node walk, and the
    # reference-order walk of every ray (UAVMVS_K7_EXACT=1) as the checker and the comparison
    m7 = min(len(pts), 1 << 18)
    p7, n7 = np.ascontiguousarray(pts[:m7]), np.ascontiguousarray(nrm[:m7])
    dirs = sv / np.linalg.norm(sv, axis=1, keepdims=True)
    rays7 = int(sum(int(((n7[i:i + 65536] @ dirs.T) >= 0.087).sum()) for i in range(0, m7, 65536)))
    res7, ms7 = {}, {}
    for mode in ("wide", "exact"):
        if mode == "exact":
            os.environ["UAVMVS_K7_EXACT"] = "1"
        c7 = api.PointCloud.create(p7, n7)
        best = None
        for _ in range(3):
            e0, e1 = ev(), ev()
            e0.record(stream)
            api.estimate_capture_difficulty(sc.max_distance, bvh, sc.num_tris, kd, c7, stream=stream)
            e1.record(stream)
            torch.cuda.synchronize()
            best = e0.elapsed_time(e1) if best is None else min(best, e0.elapsed_time(e1))
        ms7[mode] = best
        res7[mode] = c7.download_attributes()
        os.environ.pop("UAVMVS_K7_EXACT", None)
        del c7
    same7 = bool(np.array_equal(res7["wide"][0].view(np.uint32), res7["exact"][0].view(np.uint32))
                 and np.array_equal(res7["wide"][1].view(np.uint32), res7["exact"][1].view(np.uint32)))
    out["k7_closest_hit_rays_per_s"] = rays7 / (ms7["wide"] / 1e3)
    out["k7"] = {"points": m7, "directions": int(S), "rays": rays7, "ms": ms7["wide"], "ms_reference_order_walk": ms7["exact"],
                 "bits_equal_to_reference_order_walk": same7,
                 "what": "kernels.cu:448-494 on the first 2^18 points of the cloud, mesh_size = all faces; best of 3 calls, CUDA events; "
                         "rays counted on the host (ctheta >= 0.087, float64 dot products: +- a few of 1e8)"}
    if not same7:
        print("bench.py: K7 wide-node walk differs from the reference-order walk", file=sys.stderr)
        sys.exit(1)
    line["extra"] = out


def per_rank_table(blk):
    return {"columns": ["ms_per_step", "K1_ms", "k_trace_ms", "points", "sm_mhz", "k_view_masks_ms", "caps_scan_ms", "k_cull_emit_q_ms",
                        "k_retire_ms", "K3_K8_ms", "rays"],
            "rows": [[round(float(x), 4) for x in r] for r in blk["per_rank"]]}


def run_b200(a, rank, local_rank, world):
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")      # keep stdout to the one JSON line
    import torch
    import torch.distributed as dist
    from uavmvs_b200 import api

    if not torch.cuda.is_available() or api.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device (the evaluation path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    api.set_device(local_rank)
    comm = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        comm = api.Comm.from_torch_distributed()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:  # noqa: BLE001
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)") if "hbm_gbs" in peaks \
        else (6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)")
    stream = torch.cuda.current_stream()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")       # > 126 MB L2

    if a.config in (3, 4):
        return run_b200_config34(a, torch, api, stream, flush, rank, world)

    blk = evaluation_block(a, torch, dist, api, comm, rank, world, stream, flush, a.steps, a.warmup)
    sblk, a5 = None, None
    if a.config == 2 and not a.no_strong:
        # config 5 at its fixed total size over the same ranks (strong scaling), a few steps, resident inputs only
        a5 = argparse.Namespace(**{**vars(a), **CONFIGS[5], "config": 5})
        sblk = evaluation_block(a5, torch, dist, api, comm, rank, world, stream, flush, a.strong_steps, 2, want_e2e=False)
    score = blk["score"]
    rays, observations = score["rays"], score["observations"]
    sec = blk["ms"] / 1e3
    e2e_sec = blk["e2e_ms"] / 1e3
    value = rays * a.steps / sec
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": blk["ms"] / a.steps, "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_dict(a, world),
            "setup": {"tree": "GPU %s (uavmvs_bvh_build_%s, %.3f s, not timed)" % (a.tree.upper(), a.tree, blk["build_s"]),
                      "sharding": ("weak: every rank its own %d-point sample of the same mesh" % a.points) if a.scaling == "weak" else
                                  (("strong: contiguous Z-order ranges of one %d-point cloud, ceil(N / G) points per rank" % a.points) if a.shard_chunk <= 0 else
                                   ("strong: interleaved %d-point Z-order chunks of one %d-point cloud, rotating owner" % (a.shard_chunk, a.points))),
                      "collective": "ncclAllGather of one 64-byte score record per rank + fold in rank order, enqueued after K9 on the "
                                    "evaluation's stream inside uavmvs_evaluate_trajectory_sharded (inside the timed region)"
                                    if world > 1 else "none (single GPU): pack + fold kernels only"},
            "rays_per_step": rays, "observations_per_step": observations, "pairs_per_step": score["pairs"],
            "pairs_per_s": score["pairs"] * a.steps / sec, "score": score["score"],
            "per_rank": per_rank_table(blk),
            "stage_ms": {"K1": float(blk["k_ms"][0]), "K2": float(blk["k_ms"][1]), "K3_K8": float(blk["k_ms"][2]),
                         "K9": float(blk["k_ms"][3]), "score_collective": blk["coll_ms"], "of": "rank 0"},
            "e2e": {"value": rays * a.steps / e2e_sec, "unit": UNIT, "ms_per_step": blk["e2e_ms"] / a.steps,
                    "h2d_bytes_per_step": blk["views_bytes"], "d2h_bytes_per_step": int(8 * blk["n_local"] + 64 + 32),
                    "call": "uavmvs_evaluate_trajectory_sharded (host views in; per-point recon + counts + score record out, "
                            "copies overlapped with K2/K3, collective inside)",
                    "host_buffers_match_device": blk["e2e_ok"]},
            "gpu_launches": int(blk["launches"]),
            "gpu_launches_note": "counted by the library (uavmvs_launch_count) over the %d timed resident steps of rank 0; "
                                 "cub scans counted as 2 kernels each" % a.steps,
            "clocks": blk["clocks"]}
    if sblk is not None:
        ssc = sblk["score"]
        ssec = sblk["ms"] / 1e3
        line["strong"] = {"config": config_dict(a5, world), "steps": a.strong_steps, "warmup": 2,
                          "ms_per_step": sblk["ms"] / a.strong_steps, "value": ssc["rays"] * a.strong_steps / ssec,
                          "unit": UNIT, "rays_per_step": ssc["rays"], "pairs_per_s": ssc["pairs"] * a.strong_steps / ssec,
                          "score": ssc["score"], "scaling": "strong", "per_rank": per_rank_table(sblk),
                          "stage_ms": {"K1": float(sblk["k_ms"][0]), "K3_K8": float(sblk["k_ms"][2]), "K9": float(sblk["k_ms"][3]),
                                       "score_collective": sblk["coll_ms"], "of": "rank 0"},
                          "tree_build_s": sblk["build_s"], "clocks": sblk["clocks"],
                          "note": "strong-scaling efficiency at N = T(N=1) / (N x T(N)) of this block across the driver's runs"}
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- rank 0: in-run parity check, roofline of the dominant kernel, CPU baseline --------------
    import oracle as orc
    orc.build()
    team = cpu_team(orc)
    sc, pts, nrm = blk["sc"], blk["pts"], blk["nrm"]
    par, probe, tree_bytes = parity_check(orc, api, sc, blk["bvh"], blk["cloud"], pts, nrm)
    line["parity_check"] = par
    line["roofline"] = roofline(blk, probe, tree_bytes, hbm_peak, peak_src)
    if sblk is not None:
        spar, _, _ = parity_check(orc, api, sblk["sc"], sblk["bvh"], sblk["cloud"], sblk["pts"], sblk["nrm"], nviews=8)
        line["strong"]["parity_check"] = spar
        par = dict(par, ok=par["ok"] and spar["ok"])
        for k in ("bvh", "cloud", "pts", "nrm", "sc"):
            sblk.pop(k, None)
    if world == 1 and not a.no_cpu_baseline:
        tree, built_by, tree_s = cpu_tree(orc, sc, team)
        nviews = calibrate_cpu_sample(orc, tree, sc, pts, nrm, a.cpu_seconds)
        vidx = pick_views(blk["V"], nviews)
        r = cpu_step(orc, tree, sc, sc.views28()[vidx], pts, nrm)
        line["cpu_baseline"] = {
            "value": r["rays"] / r["seconds"], "unit": UNIT, "cores": team, "kind": "port",
            "pairs_per_s": r["pairs"] / r["seconds"],
            "sample": "%d of %d views (evenly strided) x all %d points, %.1f s; tree: %s (%.2f s, not timed); kernel "
                      "math: oracle port of libs/eval/kernels.cu, OpenMP over points, team of %d threads" %
                      (len(vidx), blk["V"], blk["n_local"], r["seconds"], built_by, tree_s, team)}
        del tree, r
        # SURVEY.md 8(d) "second side reference": the reference's own kernels (libs/eval/kernels.cu + libs/cacc,
        # compiled for sm_100a with its own flags as oracle/_ref/libref_cuda_fast.so) driven through its per-view
        # launch loop on the same GPU, same scene, all views.  Reported beside the CPU baseline; not the headline.
        try:
            if orc.refcuda("fast") is not None:
                t0 = time.time()
                refc = orc.RefCuda(sc.verts, sc.faces, pts, nrm, kind="fast")
                setup_s = time.time() - t0
                views_all = sc.views28()
                ref_ms = min(refc.time_evaluate(views_all, sc.max_distance, sc.width, sc.height, sc.target_recon)
                             for _ in range(3))
                ref_rows = refc.download()[0]
                line["reference_cuda"] = {
                    "value": rays / (ref_ms / 1e3), "unit": UNIT, "ms_per_step": ref_ms,
                    "kind": "reference kernels (update_observation_rays per view, process, evaluate, func), nvcc "
                            "--ftz=true --prec-div=false --prec-sqrt=false for sm_100a, acc::BVHTree built on the host "
                            "(%.2f s incl. upload, not timed)" % setup_s,
                    "observations": int(np.minimum(ref_rows, 0xFFFFFFFF).sum()),
                    "note": "best of 3, CUDA events; rays counted by the product on the same scene"}
                del refc
        except Exception as e:                      # the side reference must never break the bench line
            line["reference_cuda"] = {"unavailable": str(e)[:200]}
    if a.config == 2 and world == 1 and not a.no_extras:
        try:
            extras_config34(a, torch, api, stream, blk, line)
        except Exception as e:                      # noqa: BLE001
            line["extra"] = {"error": repr(e)[:300]}
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not par["ok"]:
        sys.stderr.write("bench.py: parity check against the oracle FAILED: %s\n" % json.dumps(par))
        sys.exit(1)


def run_b200_config34(a, torch, api, stream, flush, rank, world):
    """--config 3 / 4 as their own bench lines (one GPU; at N > 1 rank 0 alone runs and reports)."""
    from uavmvs_b200 import apps
    if rank != 0:
        return
    sc0 = make_scene(a)
    pts, nrm = rank_cloud(a, sc0, 0, 1)
    rng = np.random.default_rng(4)
    sc = type(sc0)(**{**sc0.__dict__, "points": pts, "normals": nrm, "values": rng.uniform(0.0, 1.2, len(pts)).astype(np.float32),
                      "qualities": rng.uniform(0.0, 1.0, len(pts)).astype(np.float32)})
    t0 = time.perf_counter()
    bvh = getattr(api.BVHTree, "build_" + a.tree)(sc.verts, sc.faces)
    api.synchronize()
    build_s = time.perf_counter() - t0
    sv, kd = apps.sphere_kdtree()
    S = sv.shape[0]
    ev = lambda: torch.cuda.Event(enable_timing=True)     # noqa: E731
    sampler = ClockSampler(visible_gpu_index(int(os.environ.get("LOCAL_RANK", "0"))))
    if a.config == 3:
        cloud = api.PointCloud.create(pts, nrm)
        loop = apps.OptimizerLoop(sc, bvh, cloud, kd, S, nelder_mead=True, min_distance=2.5)
        loop.populate(stream)
        loop.relax_independence()
        for _ in range(a.warmup):
            loop.iteration_nm(stream)
        torch.cuda.synchronize()
        api.launch_count(reset=True); api.k1_counters()
        sampler.start()
        ms = 0.0
        for _ in range(a.steps):
            flush.fill_(1)
            e0, e1 = ev(), ev()
            e0.record(stream)
            r = loop.iteration_nm(stream)
            e1.record(stream); e1.synchronize()
            ms += e0.elapsed_time(e1)
        rays = api.k1_counters()[3]
        extra = {"ms_per_iteration": ms / a.steps, "views_per_iteration": r["views"], "candidates_per_iteration": r["candidates"],
                 "independence": loop.independence, "score": r["score"]}
        h2d = int(r["views"] * 28 * 4 * 2 + r["candidates"] * 12); d2h = int(r["candidates"] * 16 + 12)
    else:
        cloud = api.PointCloud.create(pts, nrm, sc.values, sc.qualities)
        vol = apps.GuidanceVolume(sc, bvh, cloud, kd, S, grid=a.grid, batch=8192)
        warm = apps.GuidanceVolume(sc, bvh, cloud, kd, S, grid=a.grid, batch=8192)
        warm.subsample(8192 * max(1, min(a.warmup, 2))); warm.run(stream)
        npos = vol.subsample(a.positions or a.steps * 8192)
        torch.cuda.synchronize()
        api.launch_count(reset=True); api.k1_counters()
        sampler.start()
        e0, e1 = ev(), ev()
        e0.record(stream)
        r = vol.run(stream)
        e1.record(stream); e1.synchronize()
        ms = e0.elapsed_time(e1)
        rays = api.k1_counters()[3]
        a.steps = max(1, (npos + 8191) // 8192)
        extra = {"us_per_position": 1e3 * ms / npos, "positions": npos, "free_space_positions": vol.free_total,
                 "crc32_min_argmin_max_argmax": r["crc32"], "positions_with_signal": int((r["max"] > 0).sum())}
        h2d = int(8192 * 12); d2h = int(8192 * 16)
    clocks = sampler.stop()
    launches = api.launch_count(reset=True)
    sec = ms / 1e3
    line = {"metric": METRIC, "value": rays / sec, "unit": UNIT, "n_gpus": 1, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_dict(a, 1),
            "setup": {"tree": "GPU %s (uavmvs_bvh_build_%s, %.3f s, not timed)" % (a.tree.upper(), a.tree, build_s)},
            "rays": int(rays), **extra,
            "e2e": {"value": rays / sec, "unit": UNIT, "ms_per_step": ms / a.steps, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "note": "these workloads are host-driven through the C ABI with host inputs (views / candidate positions) and host "
                            "results (scores, arg-min / arg-max) every step: the timed region IS the end-to-end call sequence"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "issue", "kernel": "k_trace", "achieved": None, "peak": None, "unit": "Gwarp-inst/s", "frac": None,
                         "traffic": None, "note": "see the config-2 line: same traversal kernel; no ncu capture of this workload"}}
    emit(line)


def guard_stdout():
    """Keep stdout to the ONE JSON line: libraries (NCCL prints its version banner to stdout) get
    stderr; the returned file object is the real stdout."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = os.fdopen(os.dup(2), "w")
    return real


OUT = None


def emit(line):
    OUT.write(json.dumps(line) + "\n")
    OUT.flush()


def main():
    global OUT
    OUT = guard_stdout()
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if a.impl == "reference":
        run_reference(a, rank, world)
        return
    run_b200(a, rank, local_rank, world)


if __name__ == "__main__":
    main()
