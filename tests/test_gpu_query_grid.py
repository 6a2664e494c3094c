"""GPU parity of the grid-accelerated ball / cylinder queries (csrc/query_grid.cu): the same lists as the
oracle AND as the linear scan (csrc/query.cu) on every input, including the ones the grid must hand back to
the linear scan (rotation matrices that are not orthonormal, non-finite radii), radii that cover the whole
cloud, and clouds with outliers / NaN / inf coordinates.  Bit-exact (indices)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def clouds(b200, b, n, kind, seed):
    f = b200.synth.tabletop_cloud if kind == "table" else b200.synth.adversarial_cloud
    return np.stack([f(n, seed + i) for i in range(b)])


class both_paths:
    """runs the body once through the grid (threshold 0) and once through the linear scan"""

    def __init__(self, b200):
        self.ops = b200.ops

    def __iter__(self):
        old = self.ops.GRID_QUERY_MIN_N
        try:
            for name, thr in (("grid", 0), ("linear", 1 << 30)):
                self.ops.GRID_QUERY_MIN_N = thr
                yield name
        finally:
            self.ops.GRID_QUERY_MIN_N = old


def ball_both(b200, oracle_mod, x, c, r, ns):
    want = oracle_mod.ball_query(c, x, r, ns)
    for name in both_paths(b200):
        got = b200.pointnet2_utils.ball_query(r, ns, dev(x), dev(c)).cpu().numpy()
        assert (got == want).all(), name
    return want


def cyl_both(b200, oracle_mod, x, c, rot, r, hmin, hmaxs, ns):
    want = np.stack([oracle_mod.cylinder_query(c, x, rot, r, hmin, hm, ns) for hm in hmaxs])
    for name in both_paths(b200):
        got = b200.pointnet2_utils.cylinder_query_multi(r, hmin, hmaxs, ns, dev(x), dev(c), dev(rot)).cpu().numpy()
        assert (got == want).all(), name
        one = b200.pointnet2_utils.cylinder_query(r, hmin, hmaxs[-1], ns, dev(x), dev(c), dev(rot)).cpu().numpy()
        assert (one == want[-1]).all(), name
    return want


@pytest.mark.parametrize("b,n,m,r,ns,kind", [
    (1, 20000, 2048, 0.04, 64, "table"), (3, 5000, 300, 0.1, 32, "table"), (1, 1024, 512, 0.2, 16, "adv"),
    (2, 40, 9, 0.3, 16, "table"), (1, 1, 3, 0.5, 4, "table"), (1, 3000, 100, 1e-4, 8, "table"),
    (1, 6000, 64, -0.05, 16, "table"), (1, 2000, 50, 0.0, 8, "adv")])
def test_ball_query_grid_equals_linear_and_oracle(b200, oracle_mod, b, n, m, r, ns, kind):
    x = clouds(b200, b, n, kind, 3 * n + m)
    c = x[:, np.arange(m) % n].copy()
    c[:, -1] += 50.0                                          # a centre far outside the grid
    c[:, 0] -= 1e-3
    ball_both(b200, oracle_mod, x, c, r, ns)


def test_ball_query_grid_every_point_a_hit_and_huge_radius(b200, oracle_mod):
    x = clouds(b200, 1, 9000, "table", 5)
    c = x[:, :40].copy()
    # radius larger than the scene: every point is a hit, the chunk table is walked in several windows
    w = ball_both(b200, oracle_mod, x, c, 10.0, 64)
    assert (w[0, :, :64] == np.arange(64)).all()
    ball_both(b200, oracle_mod, x, c, 0.25, 128)
    for r in (float("inf"), float("nan"), 1e38):
        ball_both(b200, oracle_mod, x, c, r, 8)


def test_queries_grid_with_outliers_nan_inf(b200, oracle_mod):
    x = clouds(b200, 2, 4000, "table", 11)
    x[0, 17] = [1e6, -1e6, 3e5]                               # stretches the bounding box: cells get large
    x[0, 99] = [np.nan, 0.1, 0.1]
    x[1, 5] = [np.inf, 0.0, -np.inf]
    x[1, 3999] = [1e35, 1e35, 1e35]                           # beyond the grid's finite range
    c = x[:, np.arange(64) * 50 % 4000].copy()
    c[0, 1] = x[0, 17]                                        # centre on the outlier: finds itself
    c[0, 2] = [np.nan, np.nan, np.nan]
    c[1, 1] = x[1, 3999]
    w = ball_both(b200, oracle_mod, x, c, 0.05, 16)
    assert (w[0, 1] == 17).all() and (w[0, 2] == 0).all()
    rot = np.stack([b200.synth.seed_rotations(64, 3 + i) for i in range(2)]).reshape(2, 64, 9)
    cyl_both(b200, oracle_mod, x, c, rot, 0.05, -0.02, [0.01, 0.04], 16)


@pytest.mark.parametrize("b,n,m,ns,kind", [(1, 20000, 1024, 64, "table"), (2, 3000, 200, 16, "adv"),
                                           (1, 40, 9, 64, "table"), (3, 7000, 77, 32, "table")])
def test_cylinder_query_grid_equals_linear_and_oracle(b200, oracle_mod, b, n, m, ns, kind):
    x = clouds(b200, b, n, kind, 7 * n)
    c = x[:, np.arange(m) % n].copy()
    rot = np.stack([b200.synth.seed_rotations(m, 100 + i) for i in range(b)]).reshape(b, m, 9)
    cyl_both(b200, oracle_mod, x, c, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], ns)


def test_cylinder_query_grid_general_matrices_and_bounds(b200, oracle_mod):
    """`rot` that is not a rotation (scaled, sheared, singular, NaN): the bounding sphere does not hold, those
    centres must be answered by the linear scan.  Also depth lists that are not ascending / empty ranges."""
    rng = np.random.default_rng(4)
    n, m = 6000, 96
    x = clouds(b200, 1, n, "table", 21)
    c = x[:, np.arange(m) * 31 % n].copy()
    rot = b200.synth.seed_rotations(m, 9).reshape(m, 3, 3).copy()
    rot[0::6] *= 0.05                                         # shrinks distances: far points become hits
    rot[1::6] = rng.standard_normal((16, 3, 3)).astype(np.float32)
    rot[2::6, :, 0] = 0.0                                     # singular: x_rot == 0 for every point
    rot[3, 1, 1] = np.nan
    rot[4::6] *= 1.0005                                       # inside the orthonormality tolerance
    rot = rot.reshape(1, m, 9).astype(np.float32)
    cyl_both(b200, oracle_mod, x, c, rot, 0.05, -0.02, [0.04, 0.01, 0.03, 0.02], 32)     # not ascending
    cyl_both(b200, oracle_mod, x, c, rot, 0.05, 0.02, [0.01, 0.5], 16)                   # hmin > hmax[0]
    cyl_both(b200, oracle_mod, x, c, rot, 0.3, -0.3, [0.3], 64)                          # most of the cloud hits
    cyl_both(b200, oracle_mod, x, c, rot, 0.05, -0.02, [float("inf")], 8)


def test_ball_query_grid_200k_points(b200, oracle_mod):
    """cfg4-sized cloud: 25 KB of bitmap per output list and centre."""
    x = clouds(b200, 1, 200000, "table", 77)
    c = x[:, ::400].copy()
    ball_both(b200, oracle_mod, x, c, 0.03, 32)
    rot = b200.synth.seed_rotations(c.shape[1], 5).reshape(1, -1, 9)
    cyl_both(b200, oracle_mod, x, c, rot, 0.05, -0.02, [0.02, 0.04], 64)


@pytest.mark.parametrize("cell", [0.004, 0.02, 0.05, 0.5, 30.0])
def test_prebuilt_grid_any_cell_size(b200, oracle_mod, cell):
    """One binning serves several queries; the cell edge only changes the speed.  0.004 makes the bounding sphere
    span more cell rows than one round of the row table holds (and the grid is capped at 64 cells per axis)."""
    ops = b200.ops
    x = clouds(b200, 2, 5000, "table", 31)
    c = x[:, ::50].copy()
    m = c.shape[1]
    rot = np.stack([b200.synth.seed_rotations(m, 3 + i) for i in range(2)]).reshape(2, m, 9)
    xd, cd, rd = dev(x), dev(c), dev(rot)
    g = ops.query_grid_build(xd, cell)
    for r, ns in ((0.04, 64), (0.1, 16), (0.002, 4)):
        got = ops.ball_query(cd, xd, r, ns, grid=g).cpu().numpy()
        assert (got == oracle_mod.ball_query(c, x, r, ns)).all()
    hm = [0.01, 0.02, 0.03, 0.04]
    got = ops.cylinder_query_multi(cd, xd, rd, 0.05, -0.02, hm, 32, grid=g).cpu().numpy()
    for h, v in enumerate(hm):
        assert (got[h] == oracle_mod.cylinder_query(c, x, rot, 0.05, -0.02, v, 32)).all()
    one = ops.cylinder_query(cd, xd, rd, 0.05, -0.02, 0.04, 32, grid=g).cpu().numpy()
    assert (one == got[3]).all()
    with pytest.raises(ValueError):
        ops.ball_query(cd, xd.clone(), 0.04, 8, grid=g)


@pytest.mark.parametrize("nt", [64, 128, 256])
def test_grid_query_every_cta_width(b200, oracle_mod, nt):
    x = clouds(b200, 1, 12000, "table", 41)
    c = x[:, ::100].copy()
    rot = b200.synth.seed_rotations(c.shape[1], 5).reshape(1, -1, 9)
    b200._abi.lib.b200_query_grid_set_threads(nt)
    try:
        ball_both(b200, oracle_mod, x, c, 0.04, 64)
        cyl_both(b200, oracle_mod, x, c, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04, 0.05, 0.06, 0.07, 0.08], 64)
    finally:
        b200._abi.lib.b200_query_grid_set_threads(0)
