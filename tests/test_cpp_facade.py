"""The C++ facade (include/uavmvs_b200/cacc.hpp): a program written like the reference's
evaluate_trajectory app compiles as plain C++11 against the header and links to the C ABI
(CPU check); on a GPU it reproduces the oracle."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "evaluate_trajectory_facade.cpp")


def _compile(tmp_path):
    from uavmvs_b200 import build
    so = build.build()
    exe = str(tmp_path / "evaluate_trajectory_facade")
    subprocess.check_call(["g++", "-std=c++11", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), SRC, "-o", exe,
                           so, "-Wl,-rpath," + os.path.dirname(so)])
    return exe


def test_facade_compiles_and_links(tmp_path):
    assert os.path.exists(_compile(tmp_path))


@pytest.mark.gpu
def test_facade_runs_like_the_reference_app(tmp_path, oracle, tiny_scene, tiny_tree):
    exe = _compile(tmp_path)
    sc = tiny_scene
    obvh, a = tiny_tree
    path = str(tmp_path / "scene.bin")
    with open(path, "wb") as f:
        np.array([a.num_nodes, a.num_tris, sc.num_points, sc.num_views, sc.width, sc.height], np.uint32).tofile(f)
        np.array([sc.max_distance, sc.target_recon], np.float32).tofile(f)
        for arr in (a.nodes, a.aabbs, a.tris, a.indices, sc.points, sc.normals, sc.views28()):
            np.ascontiguousarray(arr).tofile(f)
    out = subprocess.run([exe, path], capture_output=True, text=True, check=True).stdout.split("\n")
    s, mn, mx = (float(x) for x in out[0].split())
    got = np.array([[float(x) for x in line.split()] for line in out[1:1 + sc.num_points]])
    obs = oracle.ObsRays(sc.num_points, 32)
    oracle.update_observation_rays(obvh, True, sc.views28(), sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs)
    oracle.process_observation_rays(obs)
    r = oracle.evaluate_observation_rays(obs)
    w = oracle.calculate_func_recons(r, sc.target_recon)
    assert np.array_equal(got[:, 0].astype(np.uint32), obs.num_rows)
    assert np.allclose(got[:, 1], r, rtol=1e-5, atol=0)
    assert abs(s - float(w.astype(np.float64).sum())) <= 1e-5 * s
    assert abs(mn - float(w.min())) <= 1e-5 * abs(mn) + 1e-6 and abs(mx - float(w.max())) <= 1e-5 * abs(mx) + 1e-6
