"""Generates tests/golden/ref_acc_closest.npz with the REFERENCE'S OWN acc::BVHTree::closest_point
(libs/acc/bvh_tree.h:492-548, compiled unmodified into oracle/_ref/libref_acc.so).  Run in the build
container (needs /root/reference):   python tests/golden/make_golden_closest.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from uavmvs_b200 import scenes  # noqa: E402


def queries(sc, seed=7):
    """Free-space points, points hovering near the surface, and mesh vertices themselves (distance 0)."""
    rng = np.random.default_rng(seed)
    lo, hi = sc.verts.min(0) - 10.0, sc.verts.max(0) + [10.0, 10.0, 60.0]
    return np.concatenate([rng.uniform(lo, hi, (6000, 3)),
                           sc.points[:3000] + rng.normal(0.0, 0.3, (3000, 3)),
                           sc.verts[:1000]]).astype(np.float32)


MAX_DISTS = (float("inf"), 5.0, 0.5)


def main():
    assert oracle.ref() is not None, "oracle/_ref/libref_acc.so missing"
    sc = scenes.tiny()
    rb = oracle.RefBVH(sc.verts, sc.faces, threads=2)
    q = queries(sc)
    out = {}
    for k, md in enumerate(MAX_DISTS):
        f, i, c = rb.closest_point(q, md)
        out["found%d" % k] = np.packbits(f)
        out["idx%d" % k] = i
        out["cp%d" % k] = c
    np.savez_compressed(os.path.join(HERE, "ref_acc_closest.npz"), **out)
    print("wrote ref_acc_closest.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
