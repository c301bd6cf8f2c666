"""Host-side callers of the histogram path, as the reference's apps drive it -- the loops around the kernels, not
the kernels: every number they produce comes from the C ABI (uavmvs_b200.api).

  GuidanceVolume     apps/generate_guidance_volume/generate_guidance_volume.cu:128-275 (BASELINE.json config 4):
                     candidate positions on the voxel grid above the height map; per batch of positions one batched
                     K4 + one batched K6 with device arg-min / arg-max (the reference launches K4 + K6 and copies a
                     128 x 45 image per voxel).
  OptimizerLoop      apps/optimize_trajectory/optimize_trajectory.cu:283-549 (config 3): select mutually independent
                     views, remove them (K1 populate = false), K2 + K3, candidate positions through batched K5 + K6 +
                     arg-min, re-add the moved views, K2 + K3 + K8 + K9.

`mark` arguments are optional callables invoked at stage boundaries (bench.py records CUDA events there).
"""
import zlib

import numpy as np

from . import api


def _noop(_name):
    pass


def sphere_kdtree(subdivisions=3):
    """The apps' histogram set-up (optimize_trajectory.cu:157-164): sphere directions + KD-tree over them."""
    sv = api.sphere_vertices(1.0, subdivisions)
    return sv, api.KDTree.build(sv)


class GuidanceVolume:
    def __init__(self, scene, bvh, cloud, kd, num_bins, grid=128, batch=8192):
        self.sc, self.bvh, self.cloud, self.kd, self.S, self.batch = scene, bvh, cloud, kd, num_bins, batch
        self.grid = grid
        self.positions, self.free_total, self.res = self._positions(scene.verts, grid)
        self.hists = api.Array.create(batch * num_bins)

    @staticmethod
    def _positions(verts, G):
        """generate_guidance_volume.cu:128-192: one resolution for the three axes, vertex height map, ground level =
        lowest occupied cell, positions in y / x / z order kept when at or above the height map (min_altitude 0)."""
        lo, hi = verts.min(0), verts.max(0)
        res = np.float32(max(hi[0] - lo[0], hi[1] - lo[1]) / G)
        hmap = np.full((G, G), -np.inf, np.float32)
        xi = np.minimum(((verts[:, 0] - lo[0]) / res).astype(np.int64), G - 1)
        yi = np.minimum(((verts[:, 1] - lo[1]) / res).astype(np.int64), G - 1)
        np.maximum.at(hmap, (yi, xi), verts[:, 2])
        ground = float(hmap[np.isfinite(hmap)].min())
        hmap = np.where(np.isfinite(hmap), hmap - ground, 0.0).astype(np.float32)
        ys, xs, zs = np.meshgrid(np.arange(G), np.arange(G), np.arange(G), indexing="ij")
        pz = (np.float32(ground) + zs.astype(np.float32) * res).astype(np.float32)
        keep = pz >= np.maximum(hmap[ys, xs], np.float32(0.0))
        px = ((xs.astype(np.float32) - res / np.float32(2.0)) * res + lo[0]).astype(np.float32)
        py = ((ys.astype(np.float32) - res / np.float32(2.0)) * res + lo[1]).astype(np.float32)
        pos = np.stack([px[keep], py[keep], pz[keep]], 1).astype(np.float32)
        return pos, int(keep.sum()), float(res)

    def subsample(self, count):
        """Evenly strided subset of the free-space positions (bounded runs)."""
        if count and count < self.positions.shape[0]:
            sel = np.linspace(0, self.positions.shape[0] - 1, count).astype(np.int64)
            self.positions = self.positions[sel]
        return self.positions.shape[0]

    def run(self, stream=None, mark=_noop, on_batch=None):
        """Every position through K4 + K6.  Returns dict(min, argmin, max, argmax, crc32)."""
        sc, P = self.sc, self.positions.shape[0]
        vmin, vmax = np.empty(P, np.float32), np.empty(P, np.float32)
        imin, imax = np.empty(P, np.uint32), np.empty(P, np.uint32)
        for b0 in range(0, P, self.batch):
            p = self.positions[b0:b0 + self.batch]
            B = p.shape[0]
            mark("k4")
            self.hists.null(stream=stream)
            api.populate_spherical_histogram_batch(p, sc.max_distance, self.bvh, self.cloud, self.kd, self.hists, stream=stream)
            mark("k6")
            r = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, self.kd, self.hists, B, stream=stream)
            mark("end")
            vmin[b0:b0 + B], imin[b0:b0 + B], vmax[b0:b0 + B], imax[b0:b0 + B] = r
            if on_batch:
                on_batch(b0 + B, P)
        crc = 0
        for a in (vmin, imin, vmax, imax):
            crc = zlib.crc32(a.tobytes(), crc)
        return {"min": vmin, "argmin": imin, "max": vmax, "argmax": imax, "crc32": "%08x" % crc}


def rotation_from_spherical(theta, phi):
    """utp::rotation_from_spherical (libs/utp/trajectory.cpp:14-39), float32, rows rx, ry, rz."""
    f = np.float32
    ct, st, cp, sp = f(np.cos(f(theta))), f(np.sin(f(theta))), f(np.cos(f(phi))), f(np.sin(f(phi)))
    rz = np.array([st * cp, st * sp, ct], f)
    rz = rz / f(np.sqrt(f(rz @ rz)))
    up = np.array([0, 0, -1], f)
    if not abs(float(up @ rz)) < 0.99:
        up = np.array([cp, sp, 0], f)
    rx = np.cross(up, rz).astype(f)
    rx = rx / f(np.sqrt(f(rx @ rx)))
    ry = np.cross(rz, rx).astype(f)
    ry = ry / f(np.sqrt(f(ry @ ry)))
    return np.stack([rx, ry, rz]).astype(f)


def initial_simplices(positions, scale, rng):
    """optimize_trajectory.cu:123-138, :210-226: a regular tetrahedron of the given scale around every view position,
    rotated by a random (theta, phi).  -> [V, 4, 3] float32."""
    t = np.float32(1.0 / np.sqrt(2.0))
    tet = np.array([[1, 0, -t], [-1, 0, -t], [0, 1, t], [0, -1, t]], np.float32)
    out = np.empty((len(positions), 4, 3), np.float32)
    for i, p in enumerate(np.asarray(positions, np.float32)):
        rot = rotation_from_spherical(rng.uniform(0.0, np.pi), 2.0 * rng.uniform(0.0, np.pi))
        out[i] = p[None, :] + (tet * np.float32(scale)) @ rot.T
    return out


class OptimizerLoop:
    """The state optimize_trajectory keeps on the device: observation rays of the whole trajectory, recons, wrecons --
    and, with nelder_mead=True, the simplices: an iteration then is remove -> K2/K3 -> one lock-step simplex-downhill
    step of the selected views on the device (uavmvs_nm_step: 6 batched objective evaluations' worth of K5 + K6, the
    decisions of libs/opti/nelder_mead.h in between) -> poses written on the device -> re-add -> K2/K3/K8/K9, with no copy
    to the host between its kernels; the score and the moved poses come back once, at its end."""

    def __init__(self, scene, bvh, cloud, kd, num_bins, views=None, candidates_per_view=7, want_views=16, seed=7,
                 full_eval=False, nelder_mead=False, min_distance=2.5):
        self.sc, self.bvh, self.cloud, self.kd, self.S = scene, bvh, cloud, kd, num_bins
        self.views = scene.views28().copy() if views is None else views.copy()
        self.V, self.N = self.views.shape[0], scene.num_points
        self.cpv, self.want, self.full_eval = candidates_per_view, want_views, full_eval
        self.obs = api.VectorArray.create(self.N, 32)
        self.recons, self.wrecons = api.Array.create(self.N), api.Array.create(self.N)
        self.gen = np.random.default_rng(seed)
        self.iters_left = np.full(self.V, 100.0)
        self.hists = api.Array.create(want_views * candidates_per_view * num_bins)
        self.independence = 1.0
        self.min_distance = float(min_distance)
        self.nm, self._d_sel = None, None
        if nelder_mead:
            self.nm = api.NelderMead.create(self.V, num_bins)
            # scale = 0.5 * (min_distance / 10) in the app (:213); trial positions a few metres apart exercise the kernels more
            self.nm.upload(initial_simplices(self.views[:, :3], 0.5 * self.min_distance, self.gen))

    def reeval(self, stream=None):
        sc = self.sc
        if self.full_eval:                                   # the reference's three launches over all points
            api.process_observation_rays(self.obs, stream=stream)
            api.evaluate_observation_rays(self.obs, self.recons, stream=stream)
            api.calculate_func_recons(self.recons, sc.target_recon, self.wrecons, stream=stream)
        else:                                                # changed 32-point blocks only
            api.reevaluate_observation_rays(self.obs, self.recons, sc.target_recon, self.wrecons, stream=stream)

    def populate(self, stream=None):
        """optimize_trajectory.cu:258-278: K1 for every view, then the first evaluation."""
        sc = self.sc
        self.obs.null(stream=stream)
        api.update_observation_rays_batch(True, self.views, sc.max_distance, sc.width, sc.height, self.bvh, self.cloud,
                                          self.obs, stream=stream)
        self.reeval(stream)
        return api.reduce(self.wrecons, stream=stream)[0] / self.N

    def select(self):
        """:286-310: V draws from the remaining-iterations distribution; a draw is kept when no kept view is closer
        than sqrt(independence * 2) * max_distance (:206-208)."""
        sc, views = self.sc, self.views
        min_sq = self.independence * 2.0 * sc.max_distance * sc.max_distance
        kept = []
        for idx in self.gen.choice(self.V, size=self.V, p=self.iters_left / self.iters_left.sum()):
            p = views[idx, :3]
            if idx not in kept and all(float(((p - views[k, :3]) ** 2).sum()) >= min_sq for k in kept):
                kept.append(int(idx))
                if len(kept) == self.want:
                    break
        return kept

    def relax_independence(self):
        """The reference's --independence option "reduce[s the] constraint": halve it until `want` views fit."""
        sel = self.select()
        while len(sel) < self.want and self.independence > 0.05:
            self.independence *= 0.5
            sel = self.select()
        return len(sel)

    def iteration_nm(self, stream=None, mark=_noop):
        """One optimiser iteration (:283-549) with the simplex-downhill step on the device."""
        sc, views = self.sc, self.views
        sel = self.select()
        if self._d_sel is None or self._d_sel.num_views != len(sel):
            self._d_sel = api.DeviceViews(views[sel], stream=stream)
        else:
            self._d_sel.update(views[sel], stream=stream)                      # H2D: the selected views
        d_sel = self._d_sel
        mark("remove")
        api.update_observation_rays_batch(False, d_sel, sc.max_distance, sc.width, sc.height, self.bvh, self.cloud, self.obs, stream=stream)
        mark("reeval")
        self.reeval(stream)
        mark("nm")
        self.nm.step(np.asarray(sel, np.uint32), self.min_distance, sc.max_distance, sc.target_recon, sc.calib[0], sc.width, sc.height,
                     self.bvh, self.cloud, self.kd, self.obs, self.recons, stream=stream)
        self.nm.apply_views(d_sel, stream=stream)
        mark("readd")
        api.update_observation_rays_batch(True, d_sel, sc.max_distance, sc.width, sc.height, self.bvh, self.cloud, self.obs, stream=stream)
        self.reeval(stream)
        score = api.reduce(self.wrecons, stream=stream)[0] / self.N         # D2H: the score ...
        views[sel] = d_sel.download(stream)                                  # ... and the moved poses, once per iteration
        mark("end")
        self.iters_left[sel] -= 1
        return {"views": len(sel), "candidates": 6 * len(sel), "score": score, "selected": sel}

    def iteration(self, stream=None, mark=_noop):
        """One optimiser iteration (:283-549) with random trial positions standing in for the simplex vertices
        (the regular-tetrahedron simplex of :210-226 plus three Nelder-Mead trial points = 7 per view)."""
        sc, views = self.sc, self.views
        sel = self.select()
        B = len(sel) * self.cpv
        cand = (views[sel, None, :3] + self.gen.normal(0.0, 2.0, (len(sel), self.cpv, 3))).reshape(-1, 3).astype(np.float32)
        mark("remove")
        api.update_observation_rays_batch(False, views[sel], sc.max_distance, sc.width, sc.height, self.bvh, self.cloud,
                                          self.obs, stream=stream)
        mark("reeval")
        self.reeval(stream)
        mark("k5")
        self.hists.null(stream=stream)
        api.populate_spherical_histogram_batch(cand, sc.max_distance, self.bvh, self.cloud, self.kd, self.hists,
                                               target_recon=sc.target_recon, obs_rays=self.obs, recons=self.recons,
                                               stream=stream)
        mark("k6")
        vmin, imin, vmax, imax = api.evaluate_spherical_histogram_batch(sc.calib[0], sc.width, sc.height, self.kd,
                                                                        self.hists, B, stream=stream)
        mark("readd")
        best = vmin.reshape(len(sel), self.cpv).argmin(1)
        moved = views[sel].copy()
        newpos = cand.reshape(len(sel), self.cpv, 3)[np.arange(len(sel)), best]
        for k in range(len(sel)):                            # position moves, orientation stays: w2c = [R | -R pos]
            R = moved[k, 3:19].reshape(4, 4)[:3, :3]
            moved[k, :3] = newpos[k]
            w2c = moved[k, 3:19].reshape(4, 4).copy()
            w2c[:3, 3] = -(R.astype(np.float64) @ newpos[k].astype(np.float64))
            moved[k, 3:19] = w2c.reshape(16)
        api.update_observation_rays_batch(True, moved, sc.max_distance, sc.width, sc.height, self.bvh, self.cloud,
                                          self.obs, stream=stream)
        self.reeval(stream)
        score = api.reduce(self.wrecons, stream=stream)[0] / self.N
        mark("end")
        views[sel] = moved
        self.iters_left[sel] -= 1
        return {"views": len(sel), "candidates": B, "score": score, "selected": sel, "candidates_xyz": cand,
                "vmin": vmin, "imin": imin}
