"""Proxy-cloud sampler (apps/generate_proxy_cloud/generate_proxy_cloud.cpp:84-184): the device version against the
oracle, which runs the reference's recipe with the C++ library's own std::mt19937 /
std::uniform_real_distribution<float>.  Integer / bit work: the samples, their normals and their order are
bit-identical."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from uavmvs_b200 import api as a
    assert a.device_count() >= 1
    return a


def _same(api, oracle, verts, faces, density):
    p_o, n_o = oracle.sample_proxy_cloud(verts, faces, density)
    p_g, n_g = api.sample_proxy_cloud(verts, faces, density)
    assert p_g.shape == p_o.shape
    assert np.array_equal(p_g.view(np.uint32), p_o.view(np.uint32))
    assert np.array_equal(n_g.view(np.uint32), n_o.view(np.uint32))
    return p_g, n_g


@pytest.mark.parametrize("scene,density", [("tiny", 0.5), ("dense", 2.0), ("tiny", 0.01)])
def test_sampler_bit_identical(api, oracle, scene, density):
    from uavmvs_b200 import scenes
    sc = getattr(scenes, scene)()
    p, n = _same(api, oracle, sc.verts, sc.faces, density)
    a, b, c = sc.verts[sc.faces[:, 0]], sc.verts[sc.faces[:, 1]], sc.verts[sc.faces[:, 2]]
    area = 0.5 * np.linalg.norm(np.cross(b - a, c - a), axis=1).astype(np.float64).sum()
    assert abs(len(p) - area * density) < 6 * np.sqrt(len(sc.faces)) + 1      # +1 with the fractional probability
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-5)


def test_sampler_large_faces_and_degenerate_input(api, oracle):
    """Two big triangles: thousands of draws per face (the generator's state is regenerated several times);
    a zero-area face contributes nothing (area 0 -> NaN normal is never stored)."""
    verts = np.array([[0, 0, 0], [100, 0, 0], [0, 100, 0], [100, 100, 5], [7, 7, 7]], np.float32)
    faces = np.array([[0, 1, 2], [1, 3, 2], [4, 4, 4]], np.uint32)
    p, n = _same(api, oracle, verts, faces, 1.5)
    assert 14000 < len(p) < 16500
    assert p[:, 0].min() >= 0 and p[:, 0].max() <= 100
    # samples feed the evaluation path unchanged
    cloud = api.PointCloud.create(p, n)
    assert cloud is not None


def test_sampler_empty_result(api, oracle):
    verts = np.array([[0, 0, 0], [1e-3, 0, 0], [0, 1e-3, 0]], np.float32)
    faces = np.array([[0, 1, 2]], np.uint32)
    p_o, _ = oracle.sample_proxy_cloud(verts, faces, 1e-3)
    p_g, n_g = api.sample_proxy_cloud(verts, faces, 1e-3)
    assert p_g.shape == p_o.shape and n_g.shape == p_o.shape
