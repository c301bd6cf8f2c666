import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle as orc
    orc.build()
    return orc


@pytest.fixture(scope="session")
def tiny_scene():
    from uavmvs_b200 import scenes
    return scenes.tiny()


@pytest.fixture(scope="session")
def tiny_tree(oracle, tiny_scene):
    """acc-topology tree of the tiny scene (oracle restatement of the SAH builder)."""
    bvh = oracle.BVH.build(tiny_scene.verts, tiny_scene.faces, threads=2)
    return bvh, bvh.export()


def make_rays(scene, n, seed=0):
    """Point->camera occlusion rays (libs/eval/kernels.cu:54-58) for random pairs."""
    rng = np.random.default_rng(seed)
    pi = rng.integers(0, scene.num_points, n)
    vi = rng.integers(0, scene.num_views, n)
    o = scene.points[pi]
    d = scene.view_pos[vi] - o
    l = np.sqrt((d * d).sum(1, dtype=np.float32)).astype(np.float32)
    d = (d / l[:, None]).astype(np.float32)
    return np.concatenate([o, d, (l * np.float32(0.001))[:, None], l[:, None]], 1).astype(np.float32)
