"""BASELINE.json's larger configurations at full size: cfg3 (batch 64) and cfg4 (200k points, 4096 seeds,
4096 grasps against the full cloud).  Where the oracle would take too long the reference's legacy kernel
(oracle/_ref) or size-independent properties stand in."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_cfg4_fps_200k_points_4096_seeds(b200, legacy):
    x = b200.synth.tabletop_cloud(200000, 4, extent=0.6)[None]
    got = b200.ops.furthest_point_sampling(dev(x), 4096).cpu().numpy()[0]
    assert got[0] == 0 and len(np.unique(got)) == 4096
    # furthest-point property: the distance of each pick to the set before it never increases
    p = x[0][got].astype(np.float64)
    d = np.full(4096, np.inf)
    run = []
    for j in range(1, 512):                                   # O(m^2) on the host: first 512 picks
        d = np.minimum(d, ((p - p[j - 1]) ** 2).sum(1))
        run.append(d[j])
    assert all(a >= b - 1e-9 for a, b in zip(run, run[1:]))
    if legacy is not None:
        want = legacy["ext"].furthest_point_sampling(dev(x), 4096).cpu().numpy()[0]
        assert (got == want).all()


def test_cfg4_collision_4096_grasps_against_200k_points(b200, oracle_mod):
    scene = b200.synth.tabletop_cloud(200000, 9, extent=0.6, dtype=np.float64)
    g = b200.synth.synthetic_grasps(scene, 4096, 2)
    T = np.ascontiguousarray(g[:, 13:16], dtype=np.float64)
    R = np.ascontiguousarray(g[:, 4:13], dtype=np.float64)
    thr = oracle_mod.collision_thresholds(g, 0.05)
    got = b200.ops.collision_counts(dev(scene), dev(T), dev(R), dev(np.ascontiguousarray(thr))).cpu().numpy()
    want = oracle_mod.collision_counts(scene, g, 0.05)
    assert (got == want).all()
    # counting is additive over any split of the scene
    half = b200.ops.collision_counts(dev(scene[:77777]), dev(T), dev(R), dev(np.ascontiguousarray(thr))).cpu().numpy() + \
        b200.ops.collision_counts(dev(scene[77777:]), dev(T), dev(R), dev(np.ascontiguousarray(thr))).cpu().numpy()
    assert (half == got).all()


def test_cfg4_nms_4096_idempotent_and_exact(b200, oracle_mod):
    scene = b200.synth.tabletop_cloud(200000, 9, extent=0.6)
    g = b200.synth.synthetic_grasps(scene[:400], 4096, 3)
    keep, nkeep = b200.ops.grasp_nms(dev(g), 0.03, math.cos(30.0 / 180.0 * math.pi))
    got = keep[: int(nkeep.item())].cpu().numpy()
    assert (got == oracle_mod.grasp_nms(g)).all()
    kept = b200.GraspGroup(g[got])
    assert len(kept.nms()) == len(kept)
    assert (np.diff(g[got, 0]) <= 0).all()                    # survivors in descending score order


def test_cfg3_batch64_equals_scene_by_scene(b200):
    from b200grasp.pipeline import GeometryScene
    big = GeometryScene(batch=64, num_point=20000, num_grasp=1024, seed=500, device="cuda")
    out = big.run_graph()
    torch.cuda.synchronize()
    for s in (0, 31, 63):
        one = GeometryScene(batch=1, num_point=20000, num_grasp=1024, seed=500 + s, device="cuda")
        for k in one.feat:
            one.feat[k].copy_(big.feat[k][s:s + 1])
        ref = one.run()
        for key in ("sa1_inds", "sa2_inds", "sa4_inds", "sa1_idx", "sa3_idx", "fp2_idx", "sa2_grouped_feat", "fp1_interp"):
            assert torch.equal(out[key][s], ref[key][0]), (s, key)
        assert torch.equal(out["crop_idx"][:, s], ref["crop_idx"][:, 0])
        assert torch.equal(out["collision_mask"][s], ref["collision_mask"][0])
        (k1, n1), (k2, n2) = out["kept"][s], ref["kept"][0]
        assert int(n1.item()) == int(n2.item()) and torch.equal(k1[: int(n1.item())], k2[: int(n2.item())])
