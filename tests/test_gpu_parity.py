"""GPU parity: every C-ABI entry point against the CPU oracle on the same seeded inputs.
Indices / masks / keep-sets bit-exact; float features within 1e-5 relative (stated per test)."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def clouds(b200, b, n, kind, seed):
    f = b200.synth.tabletop_cloud if kind == "table" else b200.synth.adversarial_cloud
    return np.stack([f(n, seed + i) for i in range(b)])


# ------------------------------------------------------------------ FPS

@pytest.mark.parametrize("b,n,m,kind", [
    (1, 20000, 2048, "table"), (2, 20000, 1024, "adv"), (1, 2048, 1024, "table"), (3, 1024, 512, "adv"),
    (2, 512, 256, "table"), (1, 300, 100, "adv"), (2, 37, 37, "adv"), (1, 1, 4, "table"), (1, 5000, 64, "adv"),
    (4, 777, 300, "adv"), (1, 9000, 500, "adv")])
def test_fps_bit_exact(b200, oracle_mod, b, n, m, kind):
    x = clouds(b200, b, n, kind, 11 * n + m)
    got = b200.pointnet2_utils.furthest_point_sample(dev(x), m).cpu().numpy()
    assert got.dtype == np.int32
    assert (got == oracle_mod.furthest_point_sampling(x, m)).all()


@pytest.mark.parametrize("c", [1, 2, 4, 8])
def test_fps_every_cluster_size_agrees(b200, oracle_mod, c):
    x = clouds(b200, 2, 6000, "adv", 5)
    want = oracle_mod.furthest_point_sampling(x, 700)
    b200._abi.lib.b200_fps_set_cluster(c)
    try:
        got = b200.ops.furthest_point_sampling(dev(x), 700).cpu().numpy()
    finally:
        b200._abi.lib.b200_fps_set_cluster(0)
    assert (got == want).all()


@pytest.mark.parametrize("kind,b,n,m,c,nt", [
    ("table", 1, 20000, 2048, 0, 0), ("table", 1, 20000, 2048, 8, 128), ("adv", 2, 20000, 1500, 16, 128),
    ("adv", 1, 5000, 5000, 8, 128), ("table", 2, 3500, 3000, 4, 128), ("adv", 1, 3100, 3200, 2, 128),
    ("table", 1, 49000, 700, 8, 256), ("adv", 1, 99000, 300, 16, 256), ("table", 3, 8192, 1000, 16, 128)])
def test_fps_several_samples_per_exchange_equals_chain_and_oracle(b200, oracle_mod, kind, b, n, m, c, nt):
    """fps_multi.cu accepts a run of samples per cluster exchange; fps.cu's chain kernel takes one per step.
    Both must give the oracle's indices -- on tie-heavy clouds (adv), with more samples than points, and for one,
    two and four candidates per lane (cluster size x CTA width)."""
    lib = b200._abi.lib
    x = clouds(b200, b, n, kind, 13 * n + m)
    want = oracle_mod.furthest_point_sampling(x, m)
    lib.b200_fps_set_cluster(c)
    lib.b200_fps_set_threads(nt)
    try:
        for multi in (1, 0):
            lib.b200_fps_set_multi(multi)
            got = b200.ops.furthest_point_sampling(dev(x), m).cpu().numpy()
            assert (got == want).all(), "multi=%d" % multi
    finally:
        lib.b200_fps_set_cluster(0)
        lib.b200_fps_set_threads(0)
        lib.b200_fps_set_multi(1)


def test_fps_several_samples_per_exchange_degenerate_clouds(b200, oracle_mod):
    """Every point skipped (|p|^2 <= 1e-3), all points identical, and a few live points among skipped ones:
    the acceptance chain has no / only zero-valued candidates."""
    rng = np.random.default_rng(2)
    n = 6000
    skipped = np.full((1, n, 3), 0.01, np.float32)
    same = np.full((1, n, 3), 0.37, np.float32)
    few = skipped.copy()
    few[0, rng.choice(n, 40, replace=False)] = rng.uniform(0.2, 0.4, (40, 3)).astype(np.float32)
    for x in (skipped, same, few):
        want = oracle_mod.furthest_point_sampling(x, 300)
        got = b200.ops.furthest_point_sampling(dev(x), 300).cpu().numpy()
        assert (got == want).all()


def test_fps_all_skipped_and_more_samples_than_points(b200, oracle_mod):
    x = np.full((1, 64, 3), 0.01, np.float32)
    assert (b200.ops.furthest_point_sampling(dev(x), 8).cpu().numpy() == 0).all()
    y = clouds(b200, 1, 50, "table", 3)
    got = b200.ops.furthest_point_sampling(dev(y), 80).cpu().numpy()
    assert (got == oracle_mod.furthest_point_sampling(y, 80)).all()


@pytest.mark.parametrize("kind,n0,n,m", [("table", 20000, 2048, 1024), ("adv", 20000, 2048, 1024), ("table", 6000, 1024, 512),
                                         ("adv", 3000, 512, 256), ("adv", 900, 300, 300), ("table", 4000, 4096, 2048)])
def test_fps_on_fps_ordered_prefix(b200, oracle_mod, kind, n0, n, m):
    """SA2..SA4 inputs: the points a previous FPS chose, in that order (backbone.py:110-119).  The
    speculative identity check must agree with the chain (and with the oracle) whether it holds or not."""
    x0 = clouds(b200, 2, max(n0, n), kind, 77)
    first = oracle_mod.furthest_point_sampling(x0, n)
    x = np.stack([x0[i][first[i]] for i in range(2)])
    x[1] = x[1][::-1]                                        # second cloud: not a prefix at all
    x = np.ascontiguousarray(x)
    want = oracle_mod.furthest_point_sampling(x, m)
    got = b200.ops.furthest_point_sampling(dev(x), m).cpu().numpy()
    assert (got == want).all()
    if kind == "table":
        assert (got[0] == np.arange(m)).all()                # no ties on a generic cloud: identity
    b200._abi.lib.b200_fps_set_speculation(0)
    try:
        chain = b200.ops.furthest_point_sampling(dev(x), m).cpu().numpy()
    finally:
        b200._abi.lib.b200_fps_set_speculation(1)
    assert (chain == want).all()


def test_fps_spec_with_external_verdicts(b200, oracle_mod):
    """The split form of the speculative path: prefix checks launched ahead (for several nested levels at once),
    FPS given this level's verdict and its ancestors'.  Cloud 0 is FPS-ordered (identity holds on every level),
    cloud 1 is FPS-ordered only after its first 600 points were reversed (level 1024->512 fails, deeper ones
    would pass on their own but their input is not a prefix any more), cloud 2 is in file order."""
    ops = b200.ops
    x0 = clouds(b200, 3, 6000, "table", 5)
    first = oracle_mod.furthest_point_sampling(x0, 1024)
    x = np.stack([x0[i][first[i]] for i in range(3)])
    x[1, :600] = x[1, :600][::-1]
    x[2] = x0[2, :1024]
    x = np.ascontiguousarray(x)
    xd = dev(x)
    sizes = [1024, 512, 256, 128]
    verdicts = [ops.fps_prefix_check(xd, sizes[l], sizes[l + 1]) for l in range(3)]
    heads = [v[:, 0].cpu().numpy() for v in verdicts]
    assert (heads[0] == [0, -1, -1]).all() and heads[1][0] == 0 and heads[2][0] == 0
    cur, cur_np = xd, x
    for l in range(3):
        got = ops.furthest_point_sampling_spec(cur, sizes[l + 1], verdicts[: l + 1])
        want = oracle_mod.furthest_point_sampling(cur_np, sizes[l + 1])
        assert (got.cpu().numpy() == want).all(), l
        assert (want[0] == np.arange(sizes[l + 1])).all()
        plain = ops.furthest_point_sampling(cur, sizes[l + 1]).cpu().numpy()
        assert (plain == want).all(), l
        cur_np = np.stack([cur_np[i][want[i]] for i in range(3)])
        cur = dev(cur_np)
    # a verdict that says "run the chain" for every cloud: still the same answer
    no = torch.full((3, 512), -1, dtype=torch.int32, device="cuda")
    got = ops.furthest_point_sampling_spec(xd, 512, [no]).cpu().numpy()
    assert (got == oracle_mod.furthest_point_sampling(x, 512)).all()
    with pytest.raises(ValueError):
        ops.furthest_point_sampling_spec(xd, 512, [])
    with pytest.raises(ValueError):
        ops.fps_prefix_check(xd, 2048, 512)


@pytest.mark.parametrize("b,n,m", [(1, 20000, 1024), (3, 777, 300), (2, 5, 9), (1, 1, 1)])
def test_gather_xyz_both_layouts(b200, b, n, m):
    x = clouds(b200, b, n, "table", n + m)
    rng = np.random.default_rng(m)
    idx = rng.integers(0, n, (b, m)).astype(np.int32)
    rows, cols = b200.ops.gather_xyz(dev(x), dev(idx))
    want = np.stack([x[i][idx[i]] for i in range(b)])
    assert rows.shape == (b, m, 3) and cols.shape == (b, 3, m)
    assert np.array_equal(rows.cpu().numpy(), want)
    assert np.array_equal(cols.cpu().numpy(), want.transpose(0, 2, 1))
    via = b200.ops.gather_points(dev(np.ascontiguousarray(x.transpose(0, 2, 1))), dev(idx)).cpu().numpy()
    assert np.array_equal(cols.cpu().numpy(), via)


# ------------------------------------------------------------------ queries

@pytest.mark.parametrize("b,n,m,r,ns,kind", [
    (1, 20000, 2048, 0.04, 64, "table"), (2, 2048, 1024, 0.1, 32, "table"), (1, 1024, 512, 0.2, 16, "adv"),
    (2, 512, 256, 0.3, 16, "table"), (1, 1000, 33, 0.02, 7, "adv"), (1, 5, 3, 0.5, 64, "table"),
    (1, 3000, 100, 1e-4, 8, "table")])
def test_ball_query_bit_exact(b200, oracle_mod, b, n, m, r, ns, kind):
    x = clouds(b200, b, n, kind, n + m)
    c = x[:, :m].copy() if m <= n else x[:, np.arange(m) % n]
    c[:, -1] += 50.0                                          # one centre with no neighbour at all
    got = b200.pointnet2_utils.ball_query(r, ns, dev(x), dev(c)).cpu().numpy()
    want = oracle_mod.ball_query(c, x, r, ns)
    assert got.dtype == np.int32 and (got == want).all()
    assert (got[:, -1] == 0).all()


@pytest.mark.parametrize("b,n,m,ns,kind", [(1, 20000, 1024, 64, "table"), (2, 3000, 200, 16, "adv"), (1, 40, 9, 64, "table")])
def test_cylinder_query_bit_exact_single_and_multi(b200, oracle_mod, b, n, m, ns, kind):
    x = clouds(b200, b, n, kind, 7 * n)
    c = x[:, np.arange(m) % n].copy()
    rot = np.stack([b200.synth.seed_rotations(m, 100 + i) for i in range(b)]).reshape(b, m, 9)
    hmaxs = [0.01, 0.02, 0.03, 0.04]
    multi = b200.pointnet2_utils.cylinder_query_multi(0.05, -0.02, hmaxs, ns, dev(x), dev(c), dev(rot)).cpu().numpy()
    for h, hm in enumerate(hmaxs):
        want = oracle_mod.cylinder_query(c, x, rot, 0.05, -0.02, hm, ns)
        got = b200.pointnet2_utils.cylinder_query(0.05, -0.02, hm, ns, dev(x), dev(c), dev(rot)).cpu().numpy()
        assert (got == want).all()
        assert (multi[h] == want).all()


# ------------------------------------------------------------------ gather / group / interpolate

@pytest.mark.parametrize("b,c,n,m,ns", [(1, 3, 20000, 2048, 64), (2, 128, 2048, 1024, 32), (1, 256, 512, 256, 16),
                                        (2, 5, 100, 7, 3), (1, 1, 9, 1, 1)])
def test_group_and_gather_exact(b200, oracle_mod, b, c, n, m, ns):
    rng = np.random.default_rng(c * n)
    f = rng.normal(size=(b, c, n)).astype(np.float32)
    idx = rng.integers(0, n, (b, m, ns)).astype(np.int32)
    got = b200.pointnet2_utils.grouping_operation(dev(f), dev(idx)).cpu().numpy()
    assert (got == oracle_mod.group_points(f, idx)).all()
    i1 = rng.integers(0, n, (b, m)).astype(np.int32)
    got = b200.pointnet2_utils.gather_operation(dev(f), dev(i1)).cpu().numpy()
    assert (got == oracle_mod.gather_points(f, i1)).all()


@pytest.mark.parametrize("b,c,n,m,ns", [(2, 128, 2048, 1024, 32), (1, 7, 60, 11, 3)])
def test_group_and_gather_backward(b200, oracle_mod, b, c, n, m, ns):
    rng = np.random.default_rng(ns)
    idx = rng.integers(0, n, (b, m, ns)).astype(np.int32)
    go = rng.normal(size=(b, c, m, ns)).astype(np.float32)
    f = dev(rng.normal(size=(b, c, n)).astype(np.float32)).requires_grad_(True)
    b200.pointnet2_utils.grouping_operation(f, dev(idx)).backward(dev(go))
    want = oracle_mod.group_points_grad(go, idx, n)
    assert np.allclose(f.grad.cpu().numpy(), want, rtol=RTOL, atol=1e-5)     # f32 atomics: order differs
    i1 = rng.integers(0, n, (b, m)).astype(np.int32)
    g1 = rng.normal(size=(b, c, m)).astype(np.float32)
    f2 = dev(rng.normal(size=(b, c, n)).astype(np.float32)).requires_grad_(True)
    b200.pointnet2_utils.gather_operation(f2, dev(i1)).backward(dev(g1))
    assert np.allclose(f2.grad.cpu().numpy(), oracle_mod.gather_points_grad(g1, i1, n), rtol=RTOL, atol=1e-5)


@pytest.mark.parametrize("b,n,m", [(1, 100, 200), (40, 1024, 70), (3, 60000, 40), (1, 5, 2500), (2, 300, 3)])
def test_three_nn_ties_every_lane_split(b200, oracle_mod, b, n, m):
    """Known points that are all the same point, or two alternating points: every candidate ties, the answer is
    decided by index order alone -- for each way the kernel splits the known points over lanes."""
    u = clouds(b200, b, n, "table", 9)
    for period in (1, 2, 5):
        base = clouds(b200, b, period, "table", 10 + period)
        k = np.ascontiguousarray(np.tile(base, (1, (m + period - 1) // period, 1))[:, :m])
        dist, idx = b200.pointnet2_utils.three_nn(dev(u), dev(k))
        d2, want_idx = oracle_mod.three_nn(u, k)
        assert (idx.cpu().numpy() == want_idx).all(), period
        assert (dist.cpu().numpy() == np.sqrt(d2)).all(), period


@pytest.mark.parametrize("b,n,m,c", [(1, 512, 256, 256), (2, 1024, 512, 256), (1, 33, 2, 4), (1, 10, 1500, 3),
                                     (33, 1024, 512, 2), (2, 70000, 100, 2)])
def test_three_nn_and_interpolate(b200, oracle_mod, b, n, m, c):
    u = clouds(b200, b, n, "adv", 3)
    k = clouds(b200, b, max(m, 1), "adv", 4)[:, :m]
    dist, idx = b200.pointnet2_utils.three_nn(dev(u), dev(k))
    d2, want_idx = oracle_mod.three_nn(u, k)
    assert (idx.cpu().numpy() == want_idx).all()
    assert (dist.cpu().numpy() == np.sqrt(d2)).all()          # sqrt is correctly rounded on both sides
    rng = np.random.default_rng(n)
    f = rng.normal(size=(b, c, m)).astype(np.float32)
    w = rng.uniform(size=(b, n, 3)).astype(np.float32)
    ft = dev(f).requires_grad_(True)
    out = b200.pointnet2_utils.three_interpolate(ft, idx, dev(w))
    assert np.allclose(out.detach().cpu().numpy(), oracle_mod.three_interpolate(f, want_idx, w), rtol=RTOL, atol=1e-6)
    go = rng.normal(size=(b, c, n)).astype(np.float32)
    out.backward(dev(go))
    # float atomics: each output sums ~3n/m terms in an arbitrary order
    atol = 1e-5 * max(1.0, (3.0 * n / m) ** 0.5)
    assert np.allclose(ft.grad.cpu().numpy(), oracle_mod.three_interpolate_grad(go, want_idx, w, m), rtol=RTOL, atol=atol)


# ------------------------------------------------------------------ knn

@pytest.mark.parametrize("b,d,nr,nq,k", [(1, 3, 300, 300, 1), (2, 3, 9000, 1024, 1), (1, 3, 50, 7, 4), (1, 5, 64, 33, 3),
                                         (1, 3, 1, 5, 1), (1, 3, 4, 6, 8)])
def test_knn_bit_exact(b200, oracle_mod, b, d, nr, nq, k):
    rng = np.random.default_rng(nr + nq)
    ref = (np.round(rng.normal(size=(b, d, nr)) * 4) / 4).astype(np.float32)      # ties
    q = (np.round(rng.normal(size=(b, d, nq)) * 4) / 4).astype(np.float32)
    got = b200.knn_modules.knn(dev(ref), dev(q), k).cpu().numpy()
    want = oracle_mod.knn(ref, q, k, fma_chain=True)
    kk = min(k, nr)
    assert got.dtype == np.int64 and (got[:, :kk] == want[:, :kk]).all()


def test_knn_golden_from_reference_cpu(b200):
    import glob, os
    for path in sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "knn_*.npz"))):
        z = np.load(path)
        got = b200.knn_modules.knn(dev(z["ref"]), dev(z["query"]), int(z["k"])).cpu().numpy()
        assert (got == z["inds"]).all(), path


# ------------------------------------------------------------------ collision + nms

def test_collision_golden_from_reference(b200):
    import glob, os
    for path in sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "collision_*.npz"))):
        z = np.load(path)
        det = b200.ModelFreeCollisionDetector.__new__(b200.ModelFreeCollisionDetector)
        det.finger_width, det.finger_length, det.voxel_size = 0.01, 0.06, float(z["voxel_size"])
        det.device = torch.device("cuda")
        det.scene_points = z["scene"]
        det._scene_dev = dev(z["scene"].astype(np.float64))
        gg = b200.GraspGroup(z["grasps"])
        mask, empty, ious = det.detect(gg, approach_dist=float(z["approach_dist"]), collision_thresh=float(z["collision_thresh"]),
                                       return_empty_grasp=True, empty_thresh=float(z["empty_thresh"]), return_ious=True)
        assert (mask == z["mask"]).all() and (empty == z["empty"]).all(), path
        for a, b in zip(ious, z["ious"]):
            assert (a == b).all(), path


@pytest.mark.parametrize("n,m", [(20000, 1024), (7000, 333), (1, 5)])
def test_collision_counts_bit_exact_at_size(b200, oracle_mod, n, m):
    scene = b200.synth.tabletop_cloud(n, 9, dtype=np.float64)
    g = b200.synth.synthetic_grasps(scene, m, 1)
    det = b200.ModelFreeCollisionDetector(scene, voxel_size=1e-6)         # effectively no down-sampling
    det.voxel_size = 0.01                                                 # IoU arithmetic at the test.py:31 value
    want_counts = oracle_mod.collision_counts(det.scene_points, g, 0.05)
    mask, ious = det.detect(b200.GraspGroup(g), approach_dist=0.05, collision_thresh=0.01, return_ious=True)
    want = oracle_mod.collision_finalize(want_counts, g, 0.01, 0.05, 0.01, return_ious=True)
    assert (mask == want[0]).all()
    for a, b in zip(ious, want[1]):
        assert (a == b).all()


@pytest.mark.parametrize("n,voxel,kind", [(20000, 0.01, "table"), (200000, 0.005, "table"), (5000, 0.05, "adv"),
                                          (1, 0.01, "table"), (777, 1e-5, "adv")])
def test_voxel_down_sample_bit_exact(b200, oracle_mod, n, voxel, kind):
    f = b200.synth.tabletop_cloud if kind == "table" else b200.synth.adversarial_cloud
    p = f(n, 21)                                              # float32, as callers pass it
    got = b200.voxel.voxel_down_sample(p, voxel)
    want = oracle_mod.voxel_down_sample(p.astype(np.float64), voxel)
    assert got.dtype == np.float64 and got.shape == want.shape and (got == want).all()
    det = b200.ModelFreeCollisionDetector(p, voxel_size=voxel)
    assert (det.scene_points == want).all()
    again = b200.voxel.voxel_down_sample(got, voxel)         # idempotent up to re-binning: never more rows
    assert len(again) <= len(got)


@pytest.mark.parametrize("m,crowd", [(1024, 60), (4096, 300), (333, 10), (1, 1), (64, 1), (65, 2)])
def test_nms_keep_set_bit_exact(b200, oracle_mod, m, crowd):
    scene = b200.synth.tabletop_cloud(3000, 2)
    g = b200.synth.synthetic_grasps(scene[:crowd], m, m)
    g[::7, 0] = g[0, 0]                                                   # exact score ties
    keep, nkeep = b200.ops.grasp_nms(dev(g), 0.03, math.cos(30.0 / 180.0 * math.pi))
    got = keep[: int(nkeep.item())].cpu().numpy()
    want = oracle_mod.grasp_nms(g)
    assert len(got) == len(want) and (got == want).all()
    kept = b200.GraspGroup(g).nms()
    assert (kept.grasp_group_array == g[want]).all()
    again = kept.nms()                                                    # idempotent
    assert len(again) == len(kept)
