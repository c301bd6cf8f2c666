"""The C++ facade (include/uavmvs_b200/cacc.hpp): a program written like the reference's
evaluate_trajectory app compiles as plain C++11 against the header and links to the C ABI
(CPU check); on a GPU it reproduces the oracle."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "evaluate_trajectory_facade.cpp")


SRC_OPT = os.path.join(ROOT, "tests", "cpp", "optimize_inner_loop_facade.cpp")


def _compile(tmp_path, src=SRC):
    from uavmvs_b200 import build
    so = build.build()
    exe = str(tmp_path / os.path.basename(src)[:-4])
    subprocess.check_call(["g++", "-std=c++11", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), src, "-o", exe,
                           so, "-Wl,-rpath," + os.path.dirname(so)])
    return exe


def test_facade_compiles_and_links(tmp_path):
    assert os.path.exists(_compile(tmp_path))
    assert os.path.exists(_compile(tmp_path, SRC_OPT))


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


@pytest.mark.gpu
def test_facade_optimizer_inner_loop(tmp_path, oracle, tiny_scene, tiny_tree):
    """The optimiser's inner loop through the C++ facade (remove views, K2, K3, batched K5 + K6 + arg-min over
    the candidates) against the oracle's per-candidate evaluation."""
    exe = _compile(tmp_path, SRC_OPT)
    sc = tiny_scene
    obvh, a = tiny_tree
    sv = oracle.sphere_vertices(1.0, 3)
    okd = oracle.KD.build(sv)
    kn, kv = okd.export()
    removed = np.array([2, 9, 20], np.uint32)
    rng = np.random.default_rng(8)
    cand = (sc.view_pos[removed][:, None, :] + rng.normal(0, 2.0, (3, 5, 3))).reshape(-1, 3).astype(np.float32)
    path = str(tmp_path / "opt.bin")
    with open(path, "wb") as f:
        np.array([a.num_nodes, a.num_tris, sc.num_points, sc.num_views, sc.width, sc.height, kn.shape[0], len(removed),
                  cand.shape[0]], np.uint32).tofile(f)
        np.array([sc.max_distance, sc.target_recon], np.float32).tofile(f)
        for arr in (a.nodes, a.aabbs, a.tris, a.indices, sc.points, sc.normals, sc.views28(), kn, kv, removed, cand):
            np.ascontiguousarray(arr).tofile(f)
    out = subprocess.run([exe, path], capture_output=True, text=True, check=True).stdout.split("\n")
    nc = cand.shape[0]
    best = np.array([[float(x) for x in line.split()] for line in out[:nc]])
    got = np.array([[float(x) for x in line.split()] for line in out[nc:nc + sc.num_points]])
    obs = oracle.ObsRays(sc.num_points, 32)
    views = sc.views28()
    oracle.update_observation_rays(obvh, True, views, sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs)
    oracle.update_observation_rays(obvh, False, views[removed], sc.max_distance, sc.width, sc.height, sc.points, sc.normals, obs)
    oracle.process_observation_rays(obs)
    r = oracle.evaluate_observation_rays(obs)
    assert np.array_equal(got[:, 0].astype(np.uint32), obs.num_rows)
    assert np.allclose(got[:, 1], r, rtol=1e-5, atol=0)
    S = sv.shape[0]
    for c in range(nc):
        h_ref, bins, deltas = oracle.populate_spherical_histogram_recon(
            obvh, okd, cand[c], sc.max_distance, sc.target_recon, sc.points, sc.normals, obs, got[:, 1].astype(np.float32), S)
        img = oracle.evaluate_spherical_histogram(sc.calib[0], sc.width, sc.height, sv, h_ref.astype(np.float32))
        scale = 16.0 * max(int((bins != 0xFFFFFFFF).sum()), 1)
        assert abs(best[c, 0] - float(img.min())) <= 1e-5 * scale + 1e-9
        assert 0 <= int(best[c, 1]) < 128 * 45


ACC_DEMO = os.path.join(ROOT, "oracle", "_ref", "acc_tree_facade")


def test_facade_accepts_the_reference_acc_trees_compile(oracle):
    """cacc::BVHTree<DEVICE>::create<uint, math::Vec3f>(acc_tree) and cacc::KDTree<3, DEVICE>::create<uint>(acc_kd)
    (libs/cacc/bvh_tree.h:190-194, kd_tree.h:150-154) through the facade: tests/cpp/acc_tree_facade.cpp includes the
    reference's own acc headers and compiles -Wall -Werror where the reference is present (oracle.build())."""
    if not os.path.isdir("/root/reference/libs/acc"):
        pytest.skip("the reference checkout is not on this machine")
    assert os.path.exists(ACC_DEMO)


@pytest.mark.gpu
def test_facade_accepts_the_reference_acc_trees_run():
    """The prebuilt demo (it travels in oracle/_ref/): acc-built trees uploaded through the original create<...>() calls;
    20 000 closest-hit rays and nearest-neighbour queries on the device equal acc's own CPU answers bit for bit."""
    if not os.path.exists(ACC_DEMO):
        pytest.skip("oracle/_ref/acc_tree_facade was not built (no reference checkout at build time)")
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "uavmvs_b200") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    out = subprocess.run([ACC_DEMO], capture_output=True, text=True, env=env)
    assert out.returncode == 0 and out.stdout.startswith("OK"), out.stdout + out.stderr
