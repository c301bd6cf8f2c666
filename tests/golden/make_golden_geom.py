"""Generates tests/golden/ref_geom.npz with the REFERENCE'S OWN CODE (oracle/_ref/libref_geom.so = libs/geom/sphere.h,
libs/util/numpy_io.h, libs/geom/volume_io.h compiled unmodified).  Run in the build container:

    python tests/golden/make_golden_geom.py

  sphere      generate_sphere_mesh(1.0f, 3u): the 1272 vertices, as the apps build them (first call of a process)
  npy_*       bytes written by save_numpy_file for fixed inputs (three float vectors; one int vector)
  vol         bytes written by save_volume<uint> for a fixed sparse volume
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402


def inputs():
    rng = np.random.default_rng(2024)
    f3 = rng.normal(0.0, 3.0, (3, 37)).astype(np.float32)
    i1 = rng.integers(-1000, 1000, (1, 11)).astype(np.int32)
    vmin = np.array([-12.345678, 0.5, 3.0], np.float32)
    vmax = np.array([100.25, 1e-7, 123456.789], np.float32)
    idx = np.array([1, 17, 22], np.uint32)
    imgs = rng.random((3, 45, 128, 1)).astype(np.float32)
    return f3, i1, (4, 3, 2, vmin, vmax, idx, imgs)


def main():
    R = oracle.refgeom()
    assert R is not None, "build oracle/_ref first (make -C oracle refgeom)"
    sv = np.zeros((4096, 3), np.float32)
    n = R.ref_sphere_vertices(1.0, 3, sv, 4096)
    assert n == 1272
    f3, i1, (w, h, d, vmin, vmax, idx, imgs) = inputs()
    tmp = tempfile.mkdtemp()
    p = os.path.join(tmp, "x").encode()
    assert R.ref_save_numpy_f32(p, f3, 3, f3.shape[1]) == 0
    npy_f = np.frombuffer(open(p, "rb").read(), np.uint8).copy()
    assert R.ref_save_numpy_i32(p, i1, 1, i1.shape[1]) == 0
    npy_i = np.frombuffer(open(p, "rb").read(), np.uint8).copy()
    assert R.ref_save_volume(p, w, h, d, vmin, vmax, len(idx), idx, 128, 45, 1, imgs) == 0
    vol = np.frombuffer(open(p, "rb").read(), np.uint8).copy()
    np.savez_compressed(os.path.join(HERE, "ref_geom.npz"), sphere=sv[:n].copy(), npy_f32x3=npy_f, npy_i32x1=npy_i, vol=vol)
    print("wrote ref_geom.npz:", n, "sphere vertices,", len(npy_f), len(npy_i), len(vol), "bytes")


if __name__ == "__main__":
    main()
