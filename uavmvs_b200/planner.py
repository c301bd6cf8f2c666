"""Greedy next-best-view planner: the host loop of apps/plan_trajectory/plan_trajectory.cu:224-352 as a caller
of the device path.  Per step the reference evaluates 27 candidate positions one after the
other (launch K5, launch K6, copy the 128 x 45 image back, host arg-max); here the 27 candidates go through ONE
batched K5 and ONE batched K6 + device arg-max, then the chosen view is added with K1 + K3 exactly as the
reference does (`:326-337`, no K2 in between).

Reproduced on purpose: the reference passes (avg_recon, max_distance) where K5
expects (max_distance, target_recon) (`:270-274` vs libs/eval/kernels.h:55-56).  `reference_argument_order=True`
keeps that; False passes the arguments in the order the kernel documents.

Also reproduced, and flagged: K5's delta `func(new) - func(old)` is never positive (a new observation cannot lower a
point's reconstructability), so K6's sums are <= 0 and the reference's `if (v > max)` arg-max from max = 0
(`:300-309`) never fires: every score is 0 and the loop leaves at `*it < 0.1f` (`:321`) on its first iteration.
`objective="reference"` keeps exactly that (the app as written plans nothing); `objective="improvement"` scores a
candidate by its most negative K6 pixel, the convention of optimize_trajectory.cu:416-428, and is what makes the
planner usable.

Arithmetic of the control loop is float32 in the reference's operation order (numpy float32 scalars); the
nearest-neighbour distance to the proxy cloud (`:253-257`, a host acc::KDTree there) is a brute-force float32
minimum here, which differs from the KD-tree's value only in rounding and matters only when min_distance > 0.
"""
import numpy as np

from . import api, formats

F = np.float32
PI = F(np.arccos(F(-1.0)))


def _norm(v):
    return F(np.sqrt(F(F(F(v[0] * v[0]) + F(v[1] * v[1])) + F(v[2] * v[2]))))


def _cross(a, b):
    return np.array([F(a[1] * b[2]) - F(a[2] * b[1]), F(a[2] * b[0]) - F(a[0] * b[2]),
                     F(a[0] * b[1]) - F(a[1] * b[0])], F)


def _normalized(v):
    return (v / _norm(v)).astype(F)


def candidate_offsets():
    """offsets[27], oweights[27]: plan_trajectory.cu:181-197."""
    offsets, weights = np.zeros((27, 3), F), np.zeros(27, F)
    for z in range(3):
        for y in range(3):
            for x in range(3):
                r = F(1.0) + F((z - 1) * F(0.25))
                phi = F((y - 1) * F(PI / F(8.0)))
                theta = F(PI / F(2.0)) + F((x - 1) * F(PI / F(8.0)))
                i = (z * 3 + y) * 3 + x
                offsets[i] = (F(F(r * np.sin(theta)) * np.cos(phi)), F(F(r * np.sin(theta)) * np.sin(phi)),
                              F(r * np.cos(theta)))
                weights[i] = F(1.0) - F(_norm(np.array([1.0, 0.0, 0.0], F) - offsets[i]) / F(10.0))
    return offsets, weights


def rotation_from_spherical(theta, phi):
    """libs/utp/trajectory.cpp:15-39 (row-major 3 x 3: rows rx, ry, rz = viewing direction)."""
    theta, phi = F(theta), F(phi)
    ct, st, cp, sp = F(np.cos(theta)), F(np.sin(theta)), F(np.cos(phi)), F(np.sin(phi))
    rz = _normalized(np.array([F(st * cp), F(st * sp), ct], F))
    up = np.array([0.0, 0.0, -1.0], F)
    stable = abs(F(F(F(up[0] * rz[0]) + F(up[1] * rz[1])) + F(up[2] * rz[2]))) < F(0.99)
    if not stable:
        up = np.array([cp, sp, 0.0], F)
    rx = _normalized(_cross(up, rz))
    ry = _normalized(_cross(rz, rx))
    return np.stack([rx, ry, rz]).astype(F)


class Planner:
    """State of one plan_trajectory run.  bvh_tree / cloud / kd_tree are api handles (proxy mesh, proxy cloud,
    1272-vertex sphere); cloud_points is the host copy of the cloud's positions (minimum-distance constraint)."""

    def __init__(self, bvh_tree, cloud, kd_tree, num_points, num_bins, cloud_points=None, min_distance=0.0,
                 max_distance=80.0, min_altitude=0.0, max_altitude=100.0, max_velocity=5.0, focal_length=0.86,
                 width=1920, height=1080, max_cameras=20, start_pos=(-10.0, 10.0, 2.0), start_vel=(0.1, 0.0, 0.0),
                 reference_argument_order=True, objective="reference"):
        self.bvh, self.cloud, self.kd = bvh_tree, cloud, kd_tree
        self.n, self.S = int(num_points), int(num_bins)
        self.points = None if cloud_points is None else np.asarray(cloud_points, F).reshape(-1, 3)
        self.min_distance, self.max_distance = F(min_distance), F(max_distance)
        self.min_altitude, self.max_altitude, self.max_velocity = F(min_altitude), F(max_altitude), F(max_velocity)
        self.flen, self.width, self.height = F(focal_length), int(width), int(height)
        self.reference_argument_order = reference_argument_order
        if objective not in ("reference", "improvement"):
            raise ValueError("objective must be 'reference' or 'improvement'")
        self.objective = objective
        self.offsets, self.oweights = candidate_offsets()
        self.obs = api.VectorArray.create(self.n, max_cameras)                 # ddir_hist, :160-161
        self.recons = api.Array.create(self.n)
        self.recons.null()                                                     # :164
        self.hists = api.Array.create(27 * self.S)
        self.calib = formats.views_from_trajectory(np.zeros((1, 3), F), np.eye(3, dtype=F)[None], [self.flen],
                                                   self.width, self.height)[0, 19:28].copy()
        self.pos = np.asarray(start_pos, F)
        self.vel = np.asarray(start_vel, F)
        self.avg_recon = F(1.0)                                                # :222
        self.trajectory = []                                                   # (pos, rot) per chosen view

    # ---- :236-262
    def candidates(self):
        """positions[27], penalties[27], admissible[27] for the current state."""
        speed = _norm(self.vel)
        rx = (self.vel / speed).astype(F)
        up = np.array([0.0, 0.0, 1.0], F)
        ry = _normalized(_cross(up, rx))
        rz = _normalized(_cross(rx, ry))
        pos = np.zeros((27, 3), F)
        pen = np.zeros(27, F)
        ok = np.zeros(27, bool)
        for j in range(27):
            o = self.offsets[j]
            rel = np.array([F(F(F(rx[k] * o[0]) + F(ry[k] * o[1])) + F(rz[k] * o[2])) for k in range(3)], F)
            pos[j] = self.pos + (rel * speed).astype(F)
            if _norm(pos[j] - self.pos) > self.max_velocity:
                continue
            if pos[j][2] < self.min_altitude or self.max_altitude < pos[j][2]:
                continue
            if self.points is not None and self.min_distance > 0:
                d = self.points - pos[j]
                nn = F(np.sqrt((d * d).sum(1, dtype=F).min()))
                if nn < F(2.0) * self.min_distance:
                    if nn < self.min_distance:
                        continue
                    pen[j] += F(1.0) - F(F(nn - self.min_distance) / self.min_distance)
            if pen[j] > F(1.0):
                continue
            ok[j] = True
        return pos, pen, ok

    # ---- :264-318, all admissible candidates at once
    def score(self, pos, pen, ok):
        """view_scores[27], rotations[27, 3, 3]."""
        scores, rots = np.zeros(27, F), np.tile(np.eye(3, dtype=F), (27, 1, 1))
        idx = np.nonzero(ok)[0]
        if len(idx) == 0:
            return scores, rots
        if self.reference_argument_order:
            k5_max_distance, k5_target = self.avg_recon, self.max_distance     # quirk 6, :270-274
        else:
            k5_max_distance, k5_target = self.max_distance, self.avg_recon
        self.hists.null()
        api.populate_spherical_histogram_batch(pos[idx], k5_max_distance, self.bvh, self.cloud, self.kd, self.hists,
                                               target_recon=k5_target, obs_rays=self.obs, recons=self.recons)
        vmin, imin, vmax, imax = api.evaluate_spherical_histogram_batch(self.calib, self.width, self.height, self.kd,
                                                                        self.hists, len(idx))
        if self.objective == "improvement":
            vmax, imax = (-vmin).astype(F), imin
        for b, j in enumerate(idx):
            mx, theta, phi = F(0.0), F(0.0), F(0.0)
            if vmax[b] > F(0.0):                                               # `if (v > max)` from max = 0, :300-309
                y, x = divmod(int(imax[b]), 128)
                mx = F(vmax[b])
                theta = F(F(x / F(128.0)) * F(2.0)) * PI
                phi = F(F(0.5) + F(F(y / F(45.0)) / F(2.0))) * PI
            scores[j] = F(F(mx * self.oweights[j]) * F(F(1.0) - pen[j]))
            rots[j] = rotation_from_spherical(theta, phi)
        return scores, rots

    # ---- :320-349
    def advance(self, j, pos, rots):
        """Adopt candidate j: K1 (populate) + K3, new average reconstructability."""
        rot = rots[j]
        self.vel = (pos[j] - self.pos).astype(F)
        self.pos = pos[j].copy()
        view = formats.views_from_trajectory(self.pos[None], rot[None], [self.flen], self.width, self.height)[0]
        api.update_observation_rays(True, view[0:3], self.max_distance, view[3:19], view[19:28], self.width,
                                    self.height, self.bvh, self.cloud, self.obs)
        api.evaluate_observation_rays(self.obs, self.recons)
        s, _, _ = api.reduce(self.recons)
        self.avg_recon = F(F(s) / F(self.n))
        self.trajectory.append((self.pos.copy(), rot.copy()))
        return self.avg_recon

    def step(self):
        """One iteration of the outer loop; False when the reference would stop (`*it < 0.1f`, :321)."""
        pos, pen, ok = self.candidates()
        scores, rots = self.score(pos, pen, ok)
        j = int(np.argmax(scores))                                             # std::max_element: first maximum
        if scores[j] < F(0.1):
            return False
        self.advance(j, pos, rots)
        return True


def plan_trajectory(bvh_tree, cloud, kd_tree, num_points, num_bins, num_views=500, **kw):
    """Run the greedy planner; returns (positions [V,3], rotations [V,3,3], focal lengths [V]) ready for
    formats.save_trajectory."""
    p = Planner(bvh_tree, cloud, kd_tree, num_points, num_bins, **kw)
    for _ in range(num_views):
        if not p.step():
            break
    V = len(p.trajectory)
    pos = np.array([t[0] for t in p.trajectory], F).reshape(V, 3)
    rot = np.array([t[1] for t in p.trajectory], F).reshape(V, 3, 3)
    return pos, rot, np.full(V, p.flen, F)
