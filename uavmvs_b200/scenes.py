"""Seeded synthetic inputs for the evaluation path (host side, numpy only).

These restate the *recipes* of the reference's input tools so the benchmark
configurations of BASELINE.json can be produced without MVE or any file IO:

  proxy cloud  : area-weighted surface samples with face normals, Z-order sorted
                 (apps/generate_proxy_cloud/generate_proxy_cloud.cpp:84-184)
  trajectory   : lawn-mower grid of nadir/oblique pinhole views
                 (apps/trajectory_tools/generate.cpp:160-243)
  camera model : calib = [[f*w,0,w/2],[0,f*w,h/2],[0,0,1]], w2c = [R | -R*pos]
                 (landscape branch of mve::CameraInfo::fill_calibration /
                 fill_world_to_cam, call sites apps/evaluate_trajectory/
                 evaluate_trajectory.cu:155-157)

Everything is generated in float64 and rounded once to float32; the float32
arrays are the inputs of both the CUDA path and the CPU oracle.
"""
import math
from dataclasses import dataclass, field

import numpy as np

WIDTH, HEIGHT, FLEN = 1920, 1080, 0.86


@dataclass
class Scene:
    name: str
    verts: np.ndarray          # [nv,3] f32
    faces: np.ndarray          # [nf,3] u32
    points: np.ndarray         # [N,3] f32
    normals: np.ndarray        # [N,3] f32
    values: np.ndarray         # [N] f32   (libs/util/cio.h:45 default 0)
    qualities: np.ndarray      # [N] f32   (libs/util/cio.h:46 default 1)
    view_pos: np.ndarray       # [V,3] f32
    w2c: np.ndarray            # [V,4,4] f32
    calib: np.ndarray          # [V,3,3] f32
    max_distance: float = 80.0
    target_recon: float = 3.0
    width: int = WIDTH
    height: int = HEIGHT
    meta: dict = field(default_factory=dict)

    @property
    def num_points(self):
        return self.points.shape[0]

    @property
    def num_views(self):
        return self.view_pos.shape[0]

    @property
    def num_tris(self):
        return self.faces.shape[0]

    def views28(self):
        V = self.num_views
        out = np.empty((V, 28), np.float32)
        out[:, 0:3] = self.view_pos
        out[:, 3:19] = self.w2c.reshape(V, 16)
        out[:, 19:28] = self.calib.reshape(V, 9)
        return out

    def shard(self, rank, world):
        """Contiguous Z-order range of the proxy cloud for one rank."""
        n = self.num_points
        per = (n + world - 1) // world
        lo, hi = min(rank * per, n), min((rank + 1) * per, n)
        return lo, hi


# ---------------------------------------------------------------- meshes
def icosphere(subdivisions):
    """Shared-vertex icosphere; 5 subdivisions = 10 242 vertices / 20 480 faces,
    the same counts as the reference's bundled res/meshes/sphere.ply."""
    t = (1.0 + math.sqrt(5.0)) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t),
         (0, -1, -t), (0, 1, -t), (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4),
         (11, 10, 2), (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8),
         (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    verts = np.array(v, np.float64)
    verts /= np.linalg.norm(verts, axis=1, keepdims=True)
    faces = np.array(f, np.int64)
    for _ in range(subdivisions):
        e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]])
        e.sort(axis=1)
        ue, inv = np.unique(e, axis=0, return_inverse=True)
        inv = inv.reshape(-1)
        mid = verts[ue[:, 0]] + verts[ue[:, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = verts.shape[0]
        verts = np.concatenate([verts, mid])
        nf = faces.shape[0]
        m01, m12, m20 = base + inv[:nf], base + inv[nf:2 * nf], base + inv[2 * nf:]
        a, b, c = faces[:, 0], faces[:, 1], faces[:, 2]
        faces = np.concatenate([np.stack([a, m01, m20], 1), np.stack([b, m12, m01], 1),
                                np.stack([c, m20, m12], 1), np.stack([m01, m12, m20], 1)])
    return verts, faces


def _grid_patch(origin, du, dv, nu, nv):
    """Tessellated parallelogram origin + s*du + t*dv, nu x nv quads, 2 triangles
    each, wound so the normal is du x dv."""
    s = np.linspace(0.0, 1.0, nu + 1)
    t = np.linspace(0.0, 1.0, nv + 1)
    S, T = np.meshgrid(s, t, indexing="ij")
    verts = origin[None, :] + S.reshape(-1, 1) * du[None, :] + T.reshape(-1, 1) * dv[None, :]
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    a = (i * (nv + 1) + j).reshape(-1)
    b = a + (nv + 1)
    c = b + 1
    d = a + 1
    faces = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)])
    return verts, faces


def _terrain(x, y, extent):
    k = 2.0 * math.pi / extent
    return 1.5 * np.sin(2.0 * k * x) * np.cos(3.0 * k * y) + 1.0 * np.sin(5.0 * k * (x + y))


def urban_mesh(target_tris, extent, seed, lot=32.0, street=10.0):
    """Ground (gently undulating height field) plus one yaw-rotated box building
    per lot, every face tessellated so the total is ~target_tris triangles."""
    rng = np.random.default_rng(seed)
    nl = max(1, int(extent // lot))
    lots = [(i, j) for i in range(nl) for j in range(nl)]
    bld = []
    for (i, j) in lots:
        if rng.random() < 0.12:       # plaza / park
            continue
        cx = (i + 0.5) * extent / nl + rng.uniform(-2.0, 2.0)
        cy = (j + 0.5) * extent / nl + rng.uniform(-2.0, 2.0)
        avail = extent / nl - street
        w = rng.uniform(0.45, 1.0) * avail
        d = rng.uniform(0.45, 1.0) * avail
        h = float(np.clip(rng.lognormal(math.log(20.0), 0.45), 6.0, 70.0))
        yaw = rng.uniform(-0.35, 0.35)
        # keep the rotated footprint inside the lot
        shrink = avail / (abs(math.cos(yaw)) * max(w, d) + abs(math.sin(yaw)) * max(w, d))
        if shrink < 1.0:
            w, d = w * shrink, d * shrink
        bld.append((cx, cy, w, d, h, yaw))
    area = extent * extent
    for (_, _, w, d, h, _) in bld:
        area += w * d + 2.0 * (w + d) * (h + 3.0)
    s = math.sqrt(2.0 * area / max(target_tris, 2))   # quad edge length

    parts_v, parts_f, off = [], [], 0

    def emit(v, f):
        nonlocal off
        parts_v.append(v)
        parts_f.append(f + off)
        off += v.shape[0]

    ng = max(1, int(round(extent / s)))
    gv, gf = _grid_patch(np.zeros(3), np.array([extent, 0, 0.0]), np.array([0, extent, 0.0]), ng, ng)
    gv[:, 2] = _terrain(gv[:, 0], gv[:, 1], extent)
    emit(gv, gf)
    for (cx, cy, w, d, h, yaw) in bld:
        c, sn = math.cos(yaw), math.sin(yaw)
        ux, uy = np.array([c, sn, 0.0]) * w, np.array([-sn, c, 0.0]) * d
        z0 = -3.0
        up = np.array([0.0, 0.0, h - z0])
        o = np.array([cx, cy, z0]) - 0.5 * ux - 0.5 * uy
        nw, nd, nh = (max(1, int(round(w / s))), max(1, int(round(d / s))), max(1, int(round((h - z0) / s))))
        emit(*_grid_patch(o + up, ux, uy, nw, nd))            # roof (normal +z)
        emit(*_grid_patch(o, ux, up, nw, nh))                 # wall -uy side
        emit(*_grid_patch(o + uy, up, ux, nh, nw))            # wall +uy side
        emit(*_grid_patch(o, up, uy, nh, nd))                 # wall -ux side
        emit(*_grid_patch(o + ux, uy, up, nd, nh))            # wall +ux side
    verts = np.concatenate(parts_v).astype(np.float32)
    faces = np.concatenate(parts_f).astype(np.uint32)
    return verts, faces, {"buildings": len(bld), "edge": s}


# ---------------------------------------------------------------- clouds
def _morton_order(p):
    lo, hi = p.min(0), p.max(0)
    q = ((p - lo) / np.maximum(hi - lo, 1e-9).max() * ((1 << 20) - 1)).astype(np.uint64)

    def spread(x):
        x = x & np.uint64(0x1FFFFF)
        x = (x | (x << np.uint64(32))) & np.uint64(0x1F00000000FFFF)
        x = (x | (x << np.uint64(16))) & np.uint64(0x1F0000FF0000FF)
        x = (x | (x << np.uint64(8))) & np.uint64(0x100F00F00F00F00F)
        x = (x | (x << np.uint64(4))) & np.uint64(0x10C30C30C30C30C3)
        x = (x | (x << np.uint64(2))) & np.uint64(0x1249249249249249)
        return x
    code = spread(q[:, 0]) | (spread(q[:, 1]) << np.uint64(1)) | (spread(q[:, 2]) << np.uint64(2))
    return np.argsort(code, kind="stable")


def sample_cloud(verts, faces, n, seed):
    """Area-weighted surface samples with face normals, Z-order sorted."""
    rng = np.random.default_rng(seed)
    v = verts.astype(np.float64)
    a, b, c = v[faces[:, 0]], v[faces[:, 1]], v[faces[:, 2]]
    nrm = np.cross(b - a, c - a)
    dbl = np.linalg.norm(nrm, axis=1)
    cdf = np.cumsum(dbl)
    pick = np.searchsorted(cdf, rng.random(n) * cdf[-1], side="right").clip(0, faces.shape[0] - 1)
    r1, r2 = np.sqrt(rng.random(n)), rng.random(n)
    w0, w1, w2 = 1.0 - r1, r1 * (1.0 - r2), r1 * r2
    p = w0[:, None] * a[pick] + w1[:, None] * b[pick] + w2[:, None] * c[pick]
    nn = nrm[pick] / np.maximum(dbl[pick], 1e-30)[:, None]
    order = _morton_order(p)
    return p[order].astype(np.float32), nn[order].astype(np.float32)


# ---------------------------------------------------------------- cameras
def make_view(pos, yaw, off_nadir, flen=FLEN, width=WIDTH, height=HEIGHT):
    """Pinhole view at `pos` looking `off_nadir` radians away from straight down
    toward heading `yaw` (camera axes: x right, y down, z forward)."""
    st, ct, sy, cy = math.sin(off_nadir), math.cos(off_nadir), math.sin(yaw), math.cos(yaw)
    z = np.array([st * cy, st * sy, -ct])
    x = np.array([sy, -cy, 0.0])
    y = np.cross(z, x)
    R = np.stack([x, y, z])
    w2c = np.eye(4)
    w2c[:3, :3] = R
    w2c[:3, 3] = -R @ np.asarray(pos, np.float64)
    ax = flen * width if width >= height else flen * height
    calib = np.array([[ax, 0.0, width * 0.5], [0.0, ax, height * 0.5], [0.0, 0.0, 1.0]])
    return w2c, calib


def lawnmower(x0, y0, x1, y1, altitude, fwd_overlap, side_overlap, off_nadir_deg, heading_deg=0.0,
              flen=FLEN, width=WIDTH, height=HEIGHT):
    """Serpentine grid of views; spacing from image footprint and overlaps."""
    fw = altitude / flen                       # footprint across track (image width)
    fh = fw * height / width                   # along track
    dy = max(fh * (1.0 - fwd_overlap), 1e-3)
    dx = max(fw * (1.0 - side_overlap), 1e-3)
    xs = np.arange(x0 + 0.5 * dx, x1, dx)
    ys = np.arange(y0 + 0.5 * dy, y1, dy)
    out = []
    for i, x in enumerate(xs):
        line = ys if i % 2 == 0 else ys[::-1]
        for y in line:
            yaw = math.radians(heading_deg) + (0.0 if i % 2 == 0 else math.pi)
            out.append(((x, y, altitude), yaw, math.radians(off_nadir_deg)))
    return out


def _pack_views(specs, flen=FLEN, width=WIDTH, height=HEIGHT):
    V = len(specs)
    pos = np.empty((V, 3), np.float32)
    w2c = np.empty((V, 4, 4), np.float32)
    calib = np.empty((V, 3, 3), np.float32)
    for k, (p, yaw, off) in enumerate(specs):
        p32 = np.asarray(p, np.float32)
        W, K = make_view(p32.astype(np.float64), yaw, off, flen, width, height)
        pos[k], w2c[k], calib[k] = p32, W.astype(np.float32), K.astype(np.float32)
    return pos, w2c, calib


def _thin(specs, n):
    if len(specs) <= n:
        return specs
    keep = np.linspace(0, len(specs) - 1, n).round().astype(int)
    return [specs[i] for i in keep]


def bundled_sphere():
    """The reference's res/meshes/sphere.ply (10 242 vertices / 20 480 faces, unit sphere) from the committed
    fixture tests/golden/ref_sphere_mesh.npz; the generated icosphere of the same counts if it is missing."""
    import os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "ref_sphere_mesh.npz")
    if os.path.exists(path):
        g = np.load(path)
        return g["vertices"].astype(np.float64), g["faces"].astype(np.int64)
    return icosphere(5)


# ---------------------------------------------------------------- configs
def config1(num_points=100_000, seed=7):
    """BASELINE.json configs[0]: icosphere (res/meshes/sphere.ply topology: 20 480
    faces) x30 standing on a 200 m ground quad; ~200-view lawn-mower trajectory at
    0 and 30 degrees off nadir; 100k proxy points; max_distance 80."""
    sv, sf = bundled_sphere()
    sv = sv * 30.0 + np.array([0.0, 0.0, 30.0])
    g = 100.0
    gv = np.array([[-g, -g, 0.0], [g, -g, 0.0], [g, g, 0.0], [-g, g, 0.0]])
    gf = np.array([[0, 1, 2], [0, 2, 3]]) + sv.shape[0]
    verts = np.concatenate([sv, gv]).astype(np.float32)
    faces = np.concatenate([sf, gf]).astype(np.uint32)
    pts, nrm = sample_cloud(verts, faces, num_points, seed)
    specs = lawnmower(-g, -g, g, g, 75.0, 0.80, 0.50, 0.0) + \
        lawnmower(-g, -g, g, g, 75.0, 0.80, 0.50, 30.0, heading_deg=90.0)
    pos, w2c, calib = _pack_views(specs)
    n = pts.shape[0]
    return Scene("config1_sphere", verts, faces, pts, nrm, np.zeros(n, np.float32), np.ones(n, np.float32),
                 pos, w2c, calib, 80.0, 3.0)


def urban(num_tris=1_000_000, num_points=1_000_000, num_views=2000, extent=256.0, seed=1,
          max_distance=80.0, name=None):
    """BASELINE.json configs[1] (and [4] at city scale): synthetic urban block."""
    verts, faces, meta = urban_mesh(num_tris, extent, seed)
    pts, nrm = sample_cloud(verts, faces, num_points, seed + 1)
    rng = np.random.default_rng(seed + 2)
    specs = []
    per = max(1, num_views // 4)
    # three altitudes of nadir/oblique lawn-mower lines plus an oblique ring
    for alt, off, hd in ((60.0, 0.0, 0.0), (75.0, 25.0, 90.0), (90.0, 35.0, 45.0)):
        sp = lawnmower(0.0, 0.0, extent, extent, alt, 0.85, 0.75, off, hd)
        specs += _thin(sp, per) if len(sp) >= per else sp
    ring = num_views - len(specs)
    for k in range(max(ring, 0)):
        ang = 2.0 * math.pi * k / max(ring, 1)
        r = 0.5 * extent * (0.55 + 0.45 * rng.random())
        p = (0.5 * extent + r * math.cos(ang), 0.5 * extent + r * math.sin(ang), 50.0 + 30.0 * rng.random())
        specs.append((p, ang + math.pi, math.radians(50.0 + 15.0 * rng.random())))  # look inward
    specs = specs[:num_views]
    pos, w2c, calib = _pack_views(specs)
    n = pts.shape[0]
    meta.update({"extent": extent})
    return Scene(name or f"urban_T{faces.shape[0]}_N{n}_V{len(specs)}", verts, faces, pts, nrm,
                 np.zeros(n, np.float32), np.ones(n, np.float32), pos, w2c, calib, max_distance, 3.0, meta=meta)


def tiny(num_points=4096, num_views=24, seed=3):
    """Small urban patch for fast parity tests."""
    return urban(num_tris=6000, num_points=num_points, num_views=num_views, extent=96.0, seed=seed,
                 name="tiny_urban")


def dense(num_points=8192, num_views=160, seed=5):
    """Small block flown densely: ~15 observations per point, some beyond the 32-row cap
    (exercises row order, the cap, the sort network and the pair sums)."""
    return urban(num_tris=8000, num_points=num_points, num_views=num_views, extent=64.0, seed=seed,
                 name="dense_urban")
