"""tests/golden/ref_sphere_mesh.npz = the vertex / face arrays of the reference's bundled proxy mesh
res/meshes/sphere.ply (BASELINE.json configs[0]: "a bundled res/meshes proxy mesh"), read with
uavmvs_b200.formats.load_ply.  Input fixture (data, not code); run in the build container:
    python tests/golden/make_golden_mesh.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from uavmvs_b200 import formats  # noqa: E402

SRC = "/root/reference/res/meshes/sphere.ply"

if __name__ == "__main__":
    d = formats.load_ply(SRC)
    assert d["vertices"].shape == (10242, 3) and d["faces"].shape == (20480, 3)
    np.savez_compressed(os.path.join(HERE, "ref_sphere_mesh.npz"), vertices=d["vertices"], faces=d["faces"])
    print("wrote ref_sphere_mesh.npz", os.path.getsize(os.path.join(HERE, "ref_sphere_mesh.npz")))
