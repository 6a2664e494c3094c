"""The reference's Python-level groupers and one SA/FP-style training step on the drop-in ops,
checked against plain PyTorch statements of the same arithmetic (fp32, tolerance 1e-5 relative)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def torch_group(feat, idx):
    """(B,C,N), (B,M,ns) -> (B,C,M,ns) with torch.gather (differentiable reference of grouping_operation)."""
    B, C, N = feat.shape
    _, M, ns = idx.shape
    flat = idx.reshape(B, 1, M * ns).expand(B, C, M * ns).long()
    return torch.gather(feat, 2, flat).reshape(B, C, M, ns)


@pytest.mark.parametrize("normalize", [False, True])
def test_query_and_group_matches_torch_statement(b200, oracle_mod, normalize):
    pu = b200.pointnet2_utils
    rng = np.random.default_rng(3)
    x = np.stack([b200.synth.tabletop_cloud(3000, 5), b200.synth.adversarial_cloud(3000, 6)])
    c = x[:, :200].copy()
    f = rng.normal(size=(2, 7, 3000)).astype(np.float32)
    grouper = pu.QueryAndGroup(0.1, 32, use_xyz=True, ret_grouped_xyz=True, normalize_xyz=normalize)
    ft = dev(f).requires_grad_(True)
    new_feat, gxyz = grouper(dev(x), dev(c), ft)
    idx = dev(oracle_mod.ball_query(c, x, 0.1, 32))
    want_xyz = torch_group(dev(x).transpose(1, 2).contiguous(), idx) - dev(c).transpose(1, 2).unsqueeze(-1)
    if normalize:
        want_xyz = want_xyz / 0.1
    fr = dev(f).requires_grad_(True)
    want = torch.cat([want_xyz, torch_group(fr, idx)], dim=1)
    assert new_feat.shape == (2, 10, 200, 32)
    assert torch.equal(gxyz, want_xyz) and torch.equal(new_feat, want)
    go = dev(rng.normal(size=tuple(new_feat.shape)).astype(np.float32))
    new_feat.backward(go)
    want.backward(go)
    assert torch.allclose(ft.grad, fr.grad, rtol=RTOL, atol=1e-5)


def test_cylinder_query_and_group_rotates_like_the_reference(b200, oracle_mod):
    pu = b200.pointnet2_utils
    x = np.stack([b200.synth.tabletop_cloud(4000, 8)])
    c = x[:, :128].copy()
    rot = b200.synth.seed_rotations(128, 4)[None]                       # (1,128,3,3)
    hm = [0.01, 0.02, 0.03, 0.04]
    multi = pu.cylinder_query_multi(0.05, -0.02, hm, 64, dev(x), dev(c), dev(rot.reshape(1, 128, 9)))
    for h, hmax in enumerate(hm):
        grouper = pu.CylinderQueryAndGroup(0.05, -0.02, hmax, 64, use_xyz=True, rotate_xyz=True)
        got = grouper(dev(x), dev(c), dev(rot))
        idx = dev(oracle_mod.cylinder_query(c, x, rot.reshape(1, 128, 9), 0.05, -0.02, hmax, 64))
        g = torch_group(dev(x).transpose(1, 2).contiguous(), idx) - dev(c).transpose(1, 2).unsqueeze(-1)
        want = torch.matmul(g.permute(0, 2, 3, 1).contiguous(), dev(rot)).permute(0, 3, 1, 2).contiguous()
        assert got.shape == (1, 3, 128, 64) and torch.equal(got, want)
        again = grouper.forward_with_idx(multi[h], dev(x), dev(c), dev(rot))
        assert torch.equal(again, got)


def test_group_all(b200):
    pu = b200.pointnet2_utils
    x = torch.randn(2, 50, 3, device="cuda")
    f = torch.randn(2, 4, 50, device="cuda")
    out, g = pu.GroupAll(use_xyz=True, ret_grouped_xyz=True)(x, None, f)
    assert out.shape == (2, 7, 1, 50) and torch.equal(g, x.transpose(1, 2).unsqueeze(2))
    assert torch.equal(out[:, 3:], f.unsqueeze(2))


def test_sa_fp_training_step_gradients(b200):
    """cfg5 in miniature: FPS -> gather -> QueryAndGroup -> shared MLP -> max-pool, then three_nn /
    three_interpolate back, a scalar loss and backward.  Same weights through (a) the drop-in ops and
    (b) torch.gather statements of the same ops: parameter and input-feature gradients must agree."""
    pu = b200.pointnet2_utils
    torch.manual_seed(0)
    B, N, M, C = 2, 2048, 256, 16
    xyz = dev(np.stack([b200.synth.tabletop_cloud(N, 30 + i) for i in range(B)]))
    feat0 = torch.randn(B, C, N, device="cuda")
    conv = torch.nn.Conv2d(3 + C, 32, 1).cuda()
    fp = torch.nn.Conv1d(32 + C, 8, 1).cuda()

    def run(use_ours):
        feat = feat0.clone().requires_grad_(True)
        inds = pu.furthest_point_sample(xyz, M)
        xyz_t = xyz.transpose(1, 2).contiguous()
        new_xyz = pu.gather_operation(xyz_t, inds).transpose(1, 2).contiguous()
        idx = pu.ball_query(0.1, 16, xyz, new_xyz)
        if use_ours:
            grouped = pu.QueryAndGroup(0.1, 16, use_xyz=True)(xyz, new_xyz, feat)
        else:
            gx = torch_group(xyz_t, idx) - new_xyz.transpose(1, 2).unsqueeze(-1)
            grouped = torch.cat([gx, torch_group(feat, idx)], dim=1)
        sa = torch.relu(conv(grouped)).max(dim=3)[0]                     # (B,32,M)
        dist, idx3 = pu.three_nn(xyz, new_xyz)
        w = 1.0 / (dist + 1e-8)
        w = w / w.sum(dim=2, keepdim=True)
        if use_ours:
            up = pu.three_interpolate(sa, idx3, w)
        else:
            up = (torch.gather(sa, 2, idx3.reshape(B, 1, N * 3).expand(B, 32, N * 3).long()).reshape(B, 32, N, 3)
                  * w.unsqueeze(1)).sum(-1)
        out = fp(torch.cat([up, feat], dim=1))
        loss = (out ** 2).mean()
        conv.zero_grad(); fp.zero_grad()
        loss.backward()
        return loss.item(), feat.grad.clone(), conv.weight.grad.clone(), fp.weight.grad.clone()

    a, b_ = run(True), run(False)
    assert abs(a[0] - b_[0]) <= 1e-5 * abs(b_[0])
    for x, y in zip(a[1:], b_[1:]):
        assert torch.allclose(x, y, rtol=1e-4, atol=1e-6)
