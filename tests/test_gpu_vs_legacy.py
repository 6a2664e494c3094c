"""GPU parity against the REFERENCE ITSELF: the legacy CUDA extensions compiled from
/root/reference by oracle/build_ref.py into oracle/_ref (they travel to the GPU box).
This is what pins the CPU oracle's rounding recipe and FPS tie rule to the reference's
own kernels.  Skipped when oracle/_ref is absent."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture()
def ext(legacy):
    if legacy is None:
        pytest.skip("oracle/_ref not built")
    return legacy["ext"]


@pytest.mark.parametrize("b,n,m,kind", [(1, 20000, 2048, "table"), (2, 20000, 1024, "adv"), (2, 2048, 1024, "adv"),
                                        (1, 300, 100, "adv"), (3, 700, 128, "adv")])
def test_fps_equals_legacy_kernel(b200, oracle_mod, ext, b, n, m, kind):
    f = b200.synth.tabletop_cloud if kind == "table" else b200.synth.adversarial_cloud
    x = np.stack([f(n, 31 * n + i) for i in range(b)])
    want = ext.furthest_point_sampling(dev(x), m).cpu().numpy()
    got = b200.ops.furthest_point_sampling(dev(x), m).cpu().numpy()
    assert (got == want).all()
    assert (oracle_mod.furthest_point_sampling(x, m) == want).all()      # pins the oracle too


@pytest.mark.parametrize("n,m,r,ns", [(20000, 2048, 0.04, 64), (2048, 1024, 0.1, 32), (1500, 200, 0.02, 16)])
def test_ball_and_cylinder_equal_legacy(b200, oracle_mod, ext, n, m, r, ns):
    x = np.stack([b200.synth.adversarial_cloud(n, 5), b200.synth.tabletop_cloud(n, 6)])
    c = x[:, :m].copy()
    want = ext.ball_query(dev(c), dev(x), r, ns).cpu().numpy()
    assert (b200.ops.ball_query(dev(c), dev(x), r, ns).cpu().numpy() == want).all()
    assert (oracle_mod.ball_query(c, x, r, ns) == want).all()
    rot = np.stack([b200.synth.seed_rotations(m, 1), b200.synth.seed_rotations(m, 2)]).reshape(2, m, 9)
    for hmax in (0.01, 0.04):
        want = ext.cylinder_query(dev(c), dev(x), dev(rot), 0.05, -0.02, hmax, ns).cpu().numpy()
        assert (b200.ops.cylinder_query(dev(c), dev(x), dev(rot), 0.05, -0.02, hmax, ns).cpu().numpy() == want).all()
        assert (oracle_mod.cylinder_query(c, x, rot, 0.05, -0.02, hmax, ns) == want).all()


def test_three_nn_group_interpolate_equal_legacy(b200, oracle_mod, ext):
    u = np.stack([b200.synth.adversarial_cloud(1024, 1), b200.synth.tabletop_cloud(1024, 2)])
    k = np.stack([b200.synth.adversarial_cloud(512, 3), b200.synth.tabletop_cloud(512, 4)])
    wd, wi = ext.three_nn(dev(u), dev(k))
    gd, gi = b200.ops.three_nn(dev(u), dev(k))
    assert (gi == wi).all() and (gd == wd).all()
    od, oi = oracle_mod.three_nn(u, k)
    assert (oi == wi.cpu().numpy()).all() and (od == wd.cpu().numpy()).all()
    rng = np.random.default_rng(0)
    f = dev(rng.normal(size=(2, 256, 512)).astype(np.float32))
    w = dev(rng.uniform(size=(2, 1024, 3)).astype(np.float32))
    a = b200.ops.three_interpolate(f, gi, w)
    b = ext.three_interpolate(f, wi, w)
    assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)
    idx = dev(rng.integers(0, 512, (2, 300, 32)).astype(np.int32))
    assert (b200.ops.group_points(f, idx) == ext.group_points(f, idx)).all()
    go = dev(rng.normal(size=(2, 256, 300, 32)).astype(np.float32))
    assert torch.allclose(b200.ops.group_points_grad(go, idx, 512), ext.group_points_grad(go, idx, 512), rtol=1e-4, atol=1e-4)
    g2 = dev(rng.normal(size=(2, 256, 1024)).astype(np.float32))
    assert torch.allclose(b200.ops.three_interpolate_grad(g2, gi, w, 512), ext.three_interpolate_grad(g2, wi, w, 512),
                          rtol=1e-4, atol=1e-4)


def test_knn_equals_legacy_cuda(b200, legacy):
    if legacy is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(1)
    for nr, nq, k in ((300, 300, 1), (5000, 1024, 1), (64, 40, 3)):
        ref = dev((np.round(rng.normal(size=(1, 3, nr)) * 8) / 8).astype(np.float32))
        q = dev((np.round(rng.normal(size=(1, 3, nq)) * 8) / 8).astype(np.float32))
        want = torch.empty(1, k, nq, dtype=torch.long, device="cuda")
        legacy["knn"].knn(ref, q, want)
        got = b200.knn_modules.knn(ref, q, k)
        assert (got == want).all()
