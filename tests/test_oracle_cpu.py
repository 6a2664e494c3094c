"""CPU tests: the oracle against the reference's golden vectors, its internal
consistency (closed-form FPS winner == literal tree simulation), host logic, and
that the C-ABI library loads and exports every symbol include/b200grasp.h declares."""
import ctypes
import glob
import math
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "collision_*.npz"))))
def test_collision_oracle_matches_reference_golden(oracle_mod, path):
    z = np.load(path)
    r = oracle_mod.collision_detect(z["scene"], z["grasps"], float(z["voxel_size"]), float(z["approach_dist"]),
                                    float(z["collision_thresh"]), True, float(z["empty_thresh"]), True)
    assert (r[0] == z["mask"]).all()
    assert (r[1] == z["empty"]).all()
    for a, b in zip(r[2], z["ious"]):
        assert a.dtype == np.float64 and (a == b).all()      # bit-exact IoUs
    only = oracle_mod.collision_detect(z["scene"], z["grasps"], float(z["voxel_size"]), float(z["approach_dist"]),
                                       float(z["collision_thresh"]))
    assert (only == z["mask"]).all()


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "knn_*.npz"))))
def test_knn_oracle_matches_reference_cpu_golden(oracle_mod, path):
    z = np.load(path)
    got = oracle_mod.knn(z["ref"], z["query"], int(z["k"]), fma_chain=False)
    assert got.dtype == np.int64 and (got == z["inds"]).all()
    assert got.min() >= 1                                      # 1-based


def test_golden_files_present():
    assert len(glob.glob(os.path.join(GOLD, "collision_*.npz"))) >= 4
    assert len(glob.glob(os.path.join(GOLD, "knn_*.npz"))) >= 5


@pytest.mark.parametrize("n,m", [(700, 200), (100, 50), (2048, 512), (513, 100), (20, 20), (1, 3), (37, 60)])
def test_fps_closed_form_equals_tree_simulation(oracle_mod, n, m):
    rng = np.random.default_rng(n * 1000 + m)
    x = (np.round(rng.normal(size=(2, n, 3)) * 3) / 3 + 1).astype(np.float32)   # mass ties
    x[:, : min(5, n)] = 0.001                                                    # skipped points
    a = oracle_mod.furthest_point_sampling(x, m)
    b = oracle_mod.furthest_point_sampling(x, m, tree=True)
    assert (a == b).all()
    assert (a[:, 0] == 0).all()


def test_fps_all_points_skipped_yields_zeros(oracle_mod):
    x = np.full((1, 64, 3), 0.01, np.float32)
    assert (oracle_mod.furthest_point_sampling(x, 8) == 0).all()
    assert (oracle_mod.furthest_point_sampling(x, 8, tree=True) == 0).all()


def test_ball_query_rules(oracle_mod):
    xyz = np.array([[[0, 0, 0], [0.05, 0, 0], [0.2, 0, 0], [0.01, 0, 0], [0.1, 0, 0]]], np.float32)
    new = np.array([[[0, 0, 0], [5, 5, 5]]], np.float32)
    idx = oracle_mod.ball_query(new, xyz, 0.1, 4)
    assert idx[0, 0].tolist() == [0, 1, 3, 0]        # index order, strict <, pad with first hit
    assert idx[0, 1].tolist() == [0, 0, 0, 0]        # no neighbour -> zeros


def test_cylinder_query_identity_rotation(oracle_mod):
    xyz = np.array([[[0.005, 0, 0], [0.02, 0, 0], [0.005, 0.2, 0], [-0.05, 0, 0], [0.009, 0.01, 0.01]]], np.float32)
    new = np.zeros((1, 1, 3), np.float32)
    rot = np.eye(3, dtype=np.float32).reshape(1, 1, 9)
    idx = oracle_mod.cylinder_query(new, xyz, rot, 0.05, -0.02, 0.01, 3)
    assert idx[0, 0].tolist() == [0, 4, 0]


def test_three_nn_and_interpolate(oracle_mod):
    rng = np.random.default_rng(3)
    u = rng.normal(size=(2, 40, 3)).astype(np.float32)
    k = rng.normal(size=(2, 17, 3)).astype(np.float32)
    d2, idx = oracle_mod.three_nn(u, k)
    full = ((u[:, :, None, :].astype(np.float64) - k[:, None, :, :]) ** 2).sum(-1)
    assert (np.argsort(full, axis=-1, kind="stable")[:, :, :3] == idx).all()
    assert np.allclose(np.sort(full, -1)[:, :, :3], d2, rtol=1e-5)
    d2b, idxb = oracle_mod.three_nn(u, k[:, :2])                # m < 3: unused slot stays (inf, 0)
    assert np.isinf(d2b[:, :, 2]).all() and (idxb[:, :, 2] == 0).all()
    f = rng.normal(size=(2, 5, 17)).astype(np.float32)
    w = rng.uniform(size=(2, 40, 3)).astype(np.float32)
    out = oracle_mod.three_interpolate(f, idx, w)
    ref = np.stack([(f[b][:, idx[b]] * w[b][None]).sum(-1) for b in range(2)])
    assert np.allclose(out, ref, rtol=1e-5, atol=1e-6)


def test_group_and_grads(oracle_mod):
    rng = np.random.default_rng(4)
    f = rng.normal(size=(2, 6, 50)).astype(np.float32)
    idx = rng.integers(0, 50, (2, 9, 4)).astype(np.int32)
    g = oracle_mod.group_points(f, idx)
    assert (g == np.stack([f[b][:, idx[b]] for b in range(2)])).all()
    go = rng.normal(size=g.shape).astype(np.float32)
    gp = oracle_mod.group_points_grad(go, idx, 50)
    ref = np.zeros((2, 6, 50))
    for b in range(2):
        for c in range(6):
            np.add.at(ref[b, c], idx[b].reshape(-1), go[b, c].reshape(-1).astype(np.float64))
    assert np.allclose(gp, ref, rtol=1e-5, atol=1e-6)
    i1 = rng.integers(0, 50, (2, 11)).astype(np.int32)
    assert (oracle_mod.gather_points(f, i1) == np.stack([f[b][:, i1[b]] for b in range(2)])).all()


def test_nms_oracle_properties(oracle_mod):
    import b200grasp
    scene = b200grasp.synth.tabletop_cloud(2000, 1)
    g = b200grasp.synth.synthetic_grasps(scene[:50], 400, 2)      # crowded translations
    keep = oracle_mod.grasp_nms(g)
    assert len(keep) < 400 and len(set(keep.tolist())) == len(keep)
    s = g[keep, 0]
    assert (np.diff(s) <= 0).all()                                # score order
    assert keep[0] == np.argmax(g[:, 0])
    again = oracle_mod.grasp_nms(g[keep])                         # idempotent
    assert (again == np.arange(len(keep))).all()
    assert len(oracle_mod.grasp_nms(g, translation_thresh=0.0)) == 400


def test_voxel_down_sample_oracle_properties(oracle_mod):
    p = np.random.default_rng(0).normal(size=(4000, 3))
    b = oracle_mod.voxel_down_sample(p, 0.25)
    assert b.dtype == np.float64 and 0 < len(b) < len(p)
    # every output row is the mean of the points of one voxel: the count-weighted mean is preserved
    origin = p.min(axis=0) - 0.125
    key = np.floor((p - origin) / 0.25).astype(np.int64)
    _, inv, cnt = np.unique(key, axis=0, return_inverse=True, return_counts=True)
    assert len(cnt) == len(b)
    assert np.allclose((b * cnt[:, None]).sum(0), p.sum(0), rtol=1e-12)
    one = oracle_mod.voxel_down_sample(p[:1], 0.01)
    assert (one == p[:1]).all()

def test_abi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b200grasp.h")).read()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 17
    lib = ctypes.CDLL(os.path.join(ROOT, "graspnet-baseline_b200", "lib", "libb200grasp.so"))
    for name in sorted(declared):
        assert hasattr(lib, name), name
    lib.b200_error_string.restype = ctypes.c_char_p
    assert lib.b200_error_string(0) == b"success"
    assert b"invalid" in lib.b200_error_string(-1)
    assert lib.b200_version() >= 100


def test_host_bindings_cover_abi():
    import b200grasp
    hdr = open(os.path.join(ROOT, "include", "b200grasp.h")).read()
    declared = set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(b200grasp._abi.EXPORTS)


def test_cpu_tensor_is_rejected_not_silently_computed():
    import torch
    import b200grasp
    with pytest.raises(RuntimeError):
        b200grasp.pointnet2_utils.furthest_point_sample(torch.zeros(1, 8, 3), 4)
    with pytest.raises(RuntimeError):
        b200grasp.ops.ball_query(torch.zeros(1, 2, 3), torch.zeros(1, 8, 3), 0.1, 4)


def test_graspgroup_container():
    import b200grasp
    g = b200grasp.synth.synthetic_grasps(b200grasp.synth.tabletop_cloud(100, 0), 10, 0)
    gg = b200grasp.GraspGroup(g.copy())
    assert len(gg) == 10 and gg.rotation_matrices.shape == (10, 3, 3)
    assert gg.translations.dtype == np.float32
    sub = gg[np.arange(10) % 2 == 0]
    assert len(sub) == 5
    gg.sort_by_score()
    assert (np.diff(gg.scores) <= 0).all()
    assert len(b200grasp.GraspGroup()) == 0
