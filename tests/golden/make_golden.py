"""Generates tests/golden/*.npz.  Run in the BUILD container (needs /root/reference):

    python tests/golden/make_golden.py

Two kinds of fixtures, kept apart on purpose:

  ref_*.npz   produced by the REFERENCE'S OWN CODE: libs/acc (SAH BVH build, ray
              intersection, KD-tree build + nearest neighbour) compiled unmodified from
              /root/reference by oracle/Makefile into oracle/_ref/libref_acc.so.  These
              pin the oracle's tree steps.
  port_*.npz  produced by oracle/liboracle.so (our restatement of libs/eval/kernels.cu,
              which exists in the reference only as CUDA and cannot run in the build
              container).  They freeze the restatement so that it cannot drift silently
              and give the GPU tests a fixture that does not need the oracle's build;
              they do NOT pin the restatement to the reference ("parity unpinned" for
              the kernel math).

Inputs are regenerated from seeds by uavmvs_b200.scenes, so only outputs are stored.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle  # noqa: E402
from uavmvs_b200 import scenes  # noqa: E402
from conftest import make_rays  # noqa: E402


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8).copy()


def golden_scene():
    return scenes.tiny()


def micro_mesh():
    """Box on a ground quad: silhouette edges exercise the 1e-3 slack zone."""
    v = np.array([[-20, -20, 0], [20, -20, 0], [20, 20, 0], [-20, 20, 0],
                  [-3, -2, 0], [3, -2, 0], [3, 2, 0], [-3, 2, 0],
                  [-3, -2, 8], [3, -2, 8], [3, 2, 8], [-3, 2, 8]], np.float32)
    f = np.array([[0, 1, 2], [0, 2, 3], [8, 9, 10], [8, 10, 11], [4, 5, 9], [4, 9, 8], [5, 6, 10], [5, 10, 9],
                  [6, 7, 11], [6, 11, 10], [7, 4, 8], [7, 8, 11]], np.uint32)
    return v, f


def edge_rays(n, seed):
    """Vertical rays aimed within +-5 mm of the box roof edges (slack-zone decisions)."""
    rng = np.random.default_rng(seed)
    x = np.where(rng.random(n) < 0.5, -3.0, 3.0) + rng.uniform(-5e-3, 5e-3, n)
    y = rng.uniform(-2.2, 2.2, n)
    o = np.stack([x, y, np.full(n, 20.0)], 1)
    d = np.tile(np.array([0.0, 0.0, -1.0]), (n, 1))
    return np.concatenate([o, d, np.full((n, 1), 0.0), np.full((n, 1), 50.0)], 1).astype(np.float32)


def main():
    R = oracle.ref()
    assert R is not None, "oracle/_ref/libref_acc.so missing: run `make -C oracle` with /root/reference present"
    sc = golden_scene()

    # ---- reference (libs/acc) outputs --------------------------------------------------
    rb = oracle.RefBVH(sc.verts, sc.faces, threads=2)
    ra = rb.export()
    rays = make_rays(sc, 20_000, seed=11)
    h, f, t = rb.intersect(rays)
    mv, mf = micro_mesh()
    mb = oracle.RefBVH(mv, mf, threads=1)
    ma = mb.export()
    er = edge_rays(4000, 5)
    eh, ef, et = mb.intersect(er)
    np.savez_compressed(
        os.path.join(HERE, "ref_acc_bvh.npz"),
        tiny_num_nodes=np.uint32(ra.num_nodes), tiny_nodes_sha=sha(ra.nodes), tiny_aabbs_sha=sha(ra.aabbs),
        tiny_tris_sha=sha(ra.tris), tiny_indices_sha=sha(ra.indices),
        tiny_hit=np.packbits(h), tiny_face=np.where(h > 0, f, 0xFFFFFFFF).astype(np.uint32),
        tiny_t=np.where(h > 0, t, 0).astype(np.float32),
        micro_nodes=ma.nodes, micro_aabbs=ma.aabbs, micro_tris=ma.tris, micro_indices=ma.indices,
        micro_hit=np.packbits(eh), micro_face=np.where(eh > 0, ef, 0xFFFFFFFF).astype(np.uint32))

    sv = oracle.sphere_vertices(1.0, 3)
    rk = R.ref_kd_build(np.ascontiguousarray(sv), sv.shape[0], 1)     # 1 thread: deterministic node ids
    n = R.ref_kd_num_nodes(rk)
    kn, kv = np.empty((n, 4), np.uint32), np.empty((n, 3), np.float32)
    R.ref_kd_export(rk, kn, kv)
    q = np.random.default_rng(3).normal(size=(5000, 3)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True).astype(np.float32)
    qi, qd = np.empty(5000, np.uint32), np.empty(5000, np.float32)
    R.ref_kd_find_nn(rk, q, 5000, qi, qd)
    R.ref_kd_destroy(rk)
    np.savez_compressed(os.path.join(HERE, "ref_acc_kd.npz"), sphere_sha=sha(sv), nodes=kn, verts_sha=sha(kv),
                        nn_idx=qi.astype(np.uint16), nn_dist=qd)

    # ---- restatement (libs/eval kernel math) outputs -----------------------------------
    ob = oracle.BVH.build(sc.verts, sc.faces, threads=2)
    obs = oracle.ObsRays(sc.num_points, 32)
    state, cnt = oracle.update_observation_rays(ob, True, sc.views28(), sc.max_distance, sc.width, sc.height,
                                                sc.points, sc.normals, obs, want_state=True, want_counters=True)
    rows_k1 = obs.num_rows.copy()
    data_k1_sha = sha(obs.data)
    oracle.process_observation_rays(obs)
    recons = oracle.evaluate_observation_rays(obs)
    wrecons = oracle.calculate_func_recons(recons, sc.target_recon)
    red = oracle.reduce(wrecons)
    kd = oracle.KD.build(sv)
    S = sv.shape[0]
    pos = np.array([48.0, 48.0, 45.0], np.float32)
    rng = np.random.default_rng(9)
    values = rng.uniform(0.0, 1.2, sc.num_points).astype(np.float32)
    quals = rng.uniform(0.0, 1.0, sc.num_points).astype(np.float32)
    h4, b4, s4 = oracle.populate_spherical_histogram(ob, kd, pos, sc.max_distance, sc.points, sc.normals, values, quals, S)
    h5, b5, d5 = oracle.populate_spherical_histogram_recon(ob, kd, pos, sc.max_distance, sc.target_recon, sc.points,
                                                           sc.normals, obs, recons, S)
    img = oracle.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, sv, h5.astype(np.float32))
    np.savez_compressed(
        os.path.join(HERE, "port_eval_tiny.npz"),
        state=np.packbits(state == 2), traced=np.packbits(state >= 1), counters=cnt,
        rows_k1=rows_k1.astype(np.uint16), data_k1_sha=data_k1_sha, rows=obs.num_rows.astype(np.uint16),
        data_sorted_sha=sha(obs.data), recons=recons, wrecons=wrecons, reduce=red,
        hist_pos=pos, values=values, qualities=quals, hist4=h4, hist5=h5, image5=img)
    for fn in sorted(os.listdir(HERE)):
        if fn.endswith(".npz"):
            print(fn, os.path.getsize(os.path.join(HERE, fn)), "bytes")


if __name__ == "__main__":
    main()
