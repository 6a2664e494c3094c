"""Device-resident decode -> collision -> NMS chain against the call sequence of the reference-facing drop-ins
(which are themselves pinned to the oracle / golden vectors): same collision mask, same kept rows (bit-exact)."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def make_grasps(b200, cloud, m, seed):
    rng = np.random.default_rng(seed)
    g = np.zeros((m, 17), np.float32)
    g[:, 0] = rng.uniform(0, 1, m)
    g[:, 1] = rng.uniform(0, 0.1, m)
    g[:, 2] = 0.02
    g[:, 3] = rng.integers(1, 5, m) * np.float32(0.01)
    g[:, 4:13] = b200.synth.seed_rotations(m, seed + 7).reshape(m, 9)
    g[:, 13:16] = cloud[rng.choice(len(cloud), m, replace=len(cloud) < m)]
    k = m // 3                                                          # crowds for NMS to thin out:
    g[:k, 13:16] = g[k:2 * k, 13:16] + rng.normal(0, 0.004, (k, 3)).astype(np.float32)
    g[:k, 4:13] = g[k:2 * k, 4:13]                                      # a third of the grasps shadow another one
    g[:, 16] = -1
    return g


@pytest.mark.parametrize("n,m,pad,voxel,ad,thr", [(20000, 1024, 0, 0.01, 0.05, 0.01), (6000, 300, 212, 0.005, 0.03, 0.05),
                                                  (3000, 64, 0, 0.01, 0.005, 0.01), (500, 5, 3, 0.01, 0.05, 0.2),
                                                  (800, 0, 4, 0.01, 0.05, 0.01)])
def test_device_chain_equals_dropin_sequence(b200, oracle_mod, n, m, pad, voxel, ad, thr):
    cloud = b200.synth.tabletop_cloud(n, 100 + n)
    g = make_grasps(b200, cloud, max(m, 1), n + m)[:m]
    # the drop-in sequence (numpy in / numpy out), as test.py:94-96 + gg.nms()
    det = b200.ModelFreeCollisionDetector(cloud, voxel_size=voxel)
    gg = b200.GraspGroup(g.copy())
    want_mask = det.detect(gg, approach_dist=ad, collision_thresh=thr) if m else np.zeros(0, bool)
    want = gg[~want_mask].nms().grasp_group_array if m else np.zeros((0, 17), np.float32)
    # and the oracle's mask for the same voxelised scene
    if m:
        scene = oracle_mod.voxel_down_sample(np.asarray(cloud, np.float64), voxel)
        assert (oracle_mod.collision_detect(scene, g, voxel, ad, thr) == want_mask).all()
    # device chain on rows padded with garbage beyond the valid count
    rows = np.concatenate([g, np.full((pad, 17), np.nan, np.float32)]) if pad else g
    rows_d = torch.from_numpy(np.ascontiguousarray(rows)).cuda() if len(rows) else torch.zeros((0, 17), device="cuda")
    nrows = torch.full((1,), m, dtype=torch.int32, device="cuda")
    if len(rows) == 0:
        return
    kept, nkept, mask = b200.grasp_postprocess_device(torch.from_numpy(cloud).cuda(), rows_d, nrows, voxel, ad, thr)
    k = int(nkept.item())
    assert (mask[:m].cpu().numpy().astype(bool) == want_mask).all()
    got = kept[:k].cpu().numpy()
    assert got.shape == want.shape and (got == want).all()
    if m:
        assert (b200.grasp_postprocess(cloud, g, voxel, ad, thr) == want).all()


def test_decode_then_postprocess_on_device(b200, oracle_mod):
    """pred_decode_device feeding the chain without a host round trip == oracle decode -> drop-in sequence."""
    rng = np.random.default_rng(5)
    ns, a, d = 1024, 12, 4
    f = np.float32
    cloud = b200.synth.tabletop_cloud(20000, 9)
    heads = [rng.standard_normal((1, 2, ns)).astype(f), (rng.random((1, a, ns, d)) * 1.2 - 0.1).astype(f),
             rng.standard_normal((1, a, ns, d)).astype(f), (rng.random((1, a, ns, d)) * 0.12 - 0.01).astype(f),
             (rng.random((1, a, ns, d)) * 0.06).astype(f), cloud[None, :ns].copy()]
    v = rng.standard_normal((1, ns, 3)).astype(f)
    v[..., 2] = -np.abs(v[..., 2]) - 0.3                        # views from above the table
    heads.append(v)
    keys = ("objectness_score", "grasp_score_pred", "grasp_angle_cls_pred", "grasp_width_pred", "grasp_tolerance_pred",
            "fp2_xyz", "grasp_top_view_xyz")
    ep = {k: torch.from_numpy(h).cuda() for k, h in zip(keys, heads)}
    rows, nrows = b200.pred_decode_device(ep)
    kept, nkept, mask = b200.grasp_postprocess_device(torch.from_numpy(cloud).cuda(), rows[0], nrows[0:1])
    m = int(nrows[0].item())
    g = rows[0, :m].cpu().numpy()                              # the decoded rows (checked against the oracle elsewhere)
    det = b200.ModelFreeCollisionDetector(cloud, voxel_size=0.01)
    gg = b200.GraspGroup(g)
    want_mask = det.detect(gg, approach_dist=0.05, collision_thresh=0.01)
    want = gg[~want_mask].nms().grasp_group_array
    assert (mask[:m].cpu().numpy().astype(bool) == want_mask).all()
    assert (kept[: int(nkept.item())].cpu().numpy() == want).all()
    assert 0 < len(want) < m


def test_nms_nan_and_duplicate_scores_keep_a_total_order(b200, oracle_mod):
    """NaN scores rank last (as -inf), equal scores by row number: the rank kernel always produces a permutation
    (with plain comparisons NaN rows collided and the sweep read unwritten order[] slots)"""
    import math
    rng = np.random.default_rng(4)
    cloud = b200.synth.tabletop_cloud(4000, 2)
    g = b200.synth.synthetic_grasps(cloud[:300], 700, 6)
    g[rng.choice(700, 60, replace=False), 0] = np.nan
    g[rng.choice(700, 200, replace=False), 0] = 0.5                   # mass ties
    g[5, 0], g[6, 0] = np.inf, -np.inf
    keep, nkeep = b200.ops.grasp_nms(torch.from_numpy(g).cuda(), 0.03, math.cos(30.0 / 180.0 * math.pi))
    got = keep[: int(nkeep.item())].cpu().numpy()
    want = oracle_mod.grasp_nms(g)
    assert (got == want).all()
    assert len(set(got.tolist())) == len(got) and got.min() >= 0 and got.max() < 700
    assert got[0] == 5                                                # +inf first; the NaN rows come after every number
    first_nan = np.flatnonzero(np.isnan(g[got, 0]))
    assert len(first_nan) == 0 or np.isnan(g[got[first_nan[0]:], 0]).all()


def test_voxel_down_sample_leaves_out_and_reports_invalid_points(b200, oracle_mod):
    """a NaN / inf point or a far outlier (beyond 2^21 voxels from the minimum corner) is excluded from the result and
    counted; the host API raises (open3d: "voxel_size is too small"); the detector raises at first use"""
    cloud = b200.synth.tabletop_cloud(5000, 8).astype(np.float64)
    clean = oracle_mod.voxel_down_sample(cloud, 0.01)
    bad = cloud.copy()
    bad[17] = np.nan
    bad[4000, 1] = np.inf
    out, nvox, nbad = b200.ops.voxel_down_sample(torch.from_numpy(bad).cuda(), 0.01, return_bad=True)
    assert int(nbad.item()) == 2
    got = out[: int(nvox.item())].cpu().numpy()
    want = oracle_mod.voxel_down_sample(bad, 0.01)
    assert got.shape == want.shape and (got == want).all()
    assert abs(len(got) - len(clean)) <= 2
    far = cloud.copy()
    far[0] = [1e6, 0.0, 0.6]                                          # 1e8 voxels away along x
    out, nvox, nbad = b200.ops.voxel_down_sample(torch.from_numpy(far).cuda(), 0.01, return_bad=True)
    assert int(nbad.item()) == 1
    assert (out[: int(nvox.item())].cpu().numpy() == oracle_mod.voxel_down_sample(far, 0.01)).all()
    with pytest.raises(RuntimeError):
        b200.voxel.voxel_down_sample(far, 0.01)
    det = b200.ModelFreeCollisionDetector(bad, voxel_size=0.01)
    with pytest.raises(RuntimeError):
        det.detect(b200.GraspGroup(b200.synth.synthetic_grasps(cloud, 16, 1)))
    assert (b200.voxel.voxel_down_sample(cloud, 0.01) == clean).all()  # clean input: unchanged behaviour


def test_nms_and_voxel_switches_follow_the_oracle(b200, oracle_mod):
    """The restated third-party pieces (grasp_nms, open3d voxel_down_sample: parity unpinned) expose their open
    choices as switches; every setting equals the oracle under the same setting, and the settings that must matter
    on this input do change the result."""
    import math
    from b200grasp import _abi
    lib = _abi.lib
    cloud = b200.synth.tabletop_cloud(4000, 2)
    g = b200.synth.synthetic_grasps(cloud[:150], 900, 6)
    g[:, 16] = np.arange(900) % 3                                     # three objects
    # exact boundary cases: grasp 1 sits EXACTLY translation_thresh from grasp 0 with the same rotation
    g[0, 0], g[1, 0] = 2.0, 1.9
    g[1, 4:13] = g[0, 4:13]
    g[0, 13:16] = [0.25, 0.0, 0.5]
    g[1, 13:16] = [0.25 + 0.03125, 0.0, 0.5]                          # 2^-5: exact in fp32 and fp64
    gd = torch.from_numpy(g).cuda()
    cthr = math.cos(30.0 / 180.0 * math.pi)
    results = {}
    try:
        for flags in (0, _abi.NMS_TRANS_INCLUSIVE, _abi.NMS_ROT_INCLUSIVE, _abi.NMS_SAME_OBJECT, _abi.NMS_FP32, _abi.NMS_ACOS,
                      _abi.NMS_FP32 | _abi.NMS_ACOS | _abi.NMS_TRANS_INCLUSIVE, 31):
            lib.b200_grasp_nms_set_options(flags)
            keep, nkeep = b200.ops.grasp_nms(gd, 0.03125, cthr)
            got = keep[: int(nkeep.item())].cpu().numpy()
            want = oracle_mod.post.grasp_nms(g, 0.03125, 30.0 / 180.0 * math.pi, flags=flags)
            assert got.shape == want.shape and (got == want).all(), flags
            results[flags] = got
    finally:
        lib.b200_grasp_nms_set_options(0)
    assert 1 in results[0] and 1 not in results[_abi.NMS_TRANS_INCLUSIVE]           # the strictness of < matters here
    assert len(results[_abi.NMS_SAME_OBJECT]) > len(results[0])                     # fewer pairs qualify
    pts = torch.from_numpy(cloud.astype(np.float64)).cuda()
    try:
        lib.b200_voxel_set_options(_abi.VOXEL_ORIGIN_AT_MIN)
        out, nvox = b200.ops.voxel_down_sample(pts, 0.01)
        shifted = out[: int(nvox.item())].cpu().numpy()
    finally:
        lib.b200_voxel_set_options(0)
    assert (shifted == oracle_mod.post.voxel_down_sample(cloud.astype(np.float64), 0.01, origin_at_min=True)).all()
    out, nvox = b200.ops.voxel_down_sample(pts, 0.01)
    default = out[: int(nvox.item())].cpu().numpy()
    assert (default == oracle_mod.post.voxel_down_sample(cloud.astype(np.float64), 0.01)).all()
    assert default.shape != shifted.shape or (default != shifted).any()


@pytest.mark.parametrize("n,voxel,special", [(1, 0.01, ""), (5, 0.01, ""), (1023, 0.01, ""), (1024, 0.01, ""),
                                             (1025, 0.01, "dup"), (20000, 0.01, ""), (20000, 0.0004, ""),
                                             (20000, 0.01, "bad"), (32768, 0.003, "dup"), (4000, 0.01, "allbad"),
                                             (20000, 5.0, ""), (20000, 0.01, "origin"), (20000, 0.0001, ""), (22528, 0.01, "bad")])
def test_voxel_single_cta_path_equals_multi_launch_path_and_oracle(b200, oracle_mod, n, voxel, special):
    """clouds of <= 32768 points are down-sampled by ONE single-CTA kernel (keys, radix sort over the bits in use,
    run heads, means); B200_VOXEL_MULTI_CTA forces the multi-launch path with the library sort.  Same rows, same
    counts, bit for bit, and equal to the oracle's restatement -- including duplicates (long runs), excluded points,
    > 8 bits per axis, > 32 bits in all or > 22528 points (the sort leaves shared memory for the workspace) and a cloud
    inside one voxel (no pass at all)"""
    from b200grasp import _abi
    lib = _abi.lib
    rng = np.random.default_rng(n)
    cloud = b200.synth.tabletop_cloud(n, 21).astype(np.float64)
    if "dup" in special:
        cloud[n // 2:] = cloud[rng.integers(0, n // 2, n - n // 2)]
    if special == "bad":
        cloud[3] = np.nan
        cloud[n - 1, 2] = -np.inf
        cloud[777] = [4e5, 0.1, 0.6]
    if special == "allbad":
        cloud[:] = np.nan
    flags = _abi.VOXEL_ORIGIN_AT_MIN if special == "origin" else 0
    pts = torch.from_numpy(cloud).cuda()
    res = []
    try:
        for extra in (0, _abi.VOXEL_MULTI_CTA):
            lib.b200_voxel_set_options(flags | extra)
            out, nvox, nbad = b200.ops.voxel_down_sample(pts, voxel, return_bad=True)
            res.append((out[: int(nvox.item())].cpu().numpy(), int(nbad.item())))
    finally:
        lib.b200_voxel_set_options(0)
    (small, bad_s), (multi, bad_m) = res
    assert bad_s == bad_m == {"bad": 3, "allbad": n}.get(special, 0)
    assert small.shape == multi.shape and (small == multi).all()
    if special != "allbad":
        want = oracle_mod.post.voxel_down_sample(cloud, voxel, origin_at_min=special == "origin")
        assert small.shape == want.shape and (small == want).all()
    else:
        assert len(small) == 0


@pytest.mark.parametrize("m,pad,flags", [(2048, 0, 0), (1500, 37, 0), (129, 0, 0), (1000, 24, "same_object"),
                                         (640, 0, "fp32"), (1, 0, 0), (0, 16, 0)])
def test_fused_chain_equals_separate_calls(b200, m, pad, flags):
    """csrc/post_chain.cu (five graph nodes) against collision_mask_device -> compact_rows -> grasp_nms_device ->
    gather_rows on the same rows: NaN scores, tied scores, padded garbage rows, the NMS switches; bit for bit"""
    cloud = b200.synth.tabletop_cloud(20000, 77)
    g = make_grasps(b200, cloud, max(m, 1), 900 + m)[:m]
    if m >= 64:
        g[5:25, 0] = np.float32(0.5)                                    # ties
        g[40:44, 0] = np.nan                                            # NaN scores rank last, in row order
        g[:, 16] = np.arange(m) % 3
    rows = np.concatenate([g, np.full((pad, 17), np.nan, np.float32)]) if pad else g
    rows_d = torch.from_numpy(np.ascontiguousarray(rows)).cuda()
    nrows = torch.full((1,), m, dtype=torch.int32, device="cuda")
    cloud_d = torch.from_numpy(cloud).cuda()
    opt = {0: 0, "same_object": b200._abi.NMS_SAME_OBJECT, "fp32": b200._abi.NMS_FP32 | b200._abi.NMS_ROT_INCLUSIVE}[flags]
    b200._abi.lib.b200_grasp_nms_set_options(opt)
    try:
        a = b200.grasp_postprocess_device(cloud_d, rows_d, nrows, 0.01, 0.05, 0.01)
        c = b200.grasp_postprocess_device(cloud_d, rows_d, nrows, 0.01, 0.05, 0.01, separate_calls=True)
    finally:
        b200._abi.lib.b200_grasp_nms_set_options(0)
    ka, kc = int(a[1].item()), int(c[1].item())
    assert ka == kc
    if m >= 64:
        assert 0 < ka < m
    assert torch.equal(a[2][:m], c[2][:m])
    ra, rc = a[0][:ka].cpu().numpy(), c[0][:kc].cpu().numpy()
    assert (ra.view(np.uint32) == rc.view(np.uint32)).all()             # bit patterns: the rows hold NaN scores


def test_fused_chain_refuses_more_rows_than_its_limit(b200):
    rows = torch.zeros((b200.ops.POST_CHAIN_MAX_M + 1, 17), device="cuda")
    scene = torch.zeros((4, 3), dtype=torch.float64, device="cuda")
    n = torch.full((1,), 4, dtype=torch.int32, device="cuda")
    with pytest.raises(ValueError):
        b200.ops.grasp_post_chain(scene, n, rows, n, 0.01, 0.05, 0.01, 0.03, 0.8)
    ws = torch.empty((1 << 20,), dtype=torch.uint8, device="cuda")
    rc = b200._abi.lib.b200_grasp_post_chain(scene.data_ptr(), n.data_ptr(), 4, rows.data_ptr(), n.data_ptr(),
                                             rows.shape[0], 0.01, 0.05, 0.01, 0.03, 0.8, ws.data_ptr(), ws.data_ptr(),
                                             n.data_ptr(), None, ws.data_ptr(), ws.numel(), None)
    assert rc == -2                                                      # B200_ECAPACITY
