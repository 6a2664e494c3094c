"""GPU parity of the fused group -> shared MLP -> max-pool kernel (csrc/mlp_fused.cu, tcgen05 tensor cores):
against the outputs of the reference's SharedMLP (tests/golden/mlp_*.npz), against the oracle on seeded inputs at
the sizes GraspNet runs (SA1, CloudCrop), and the drop-in modules against their own composed (unfused) path.
Tolerance: 1e-5 relative to the largest feature (north_star) for the fp32-grade mode, 2e-3 for plain TF32."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
RTOL = 1e-5


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def device_layers(b200, layers):
    (w1, b1), (w2, b2), (w3, b3) = layers
    return [(dev(w1), dev(b1)), (b200.ops.mlp_pack_weights(dev(w2)), dev(b2), w2.shape[0]),
            (b200.ops.mlp_pack_weights(dev(w3)), dev(b3), w3.shape[0])]


def close(got, want, rtol):
    return np.allclose(got, want, rtol=rtol, atol=rtol * np.abs(want).max())


def feat_layers(b200, layers):
    """first layer over [xyz | features] -> packed over [features | xyz | zero pad] (include/b200grasp.h)"""
    (w1, b1), (w2, b2), (w3, b3) = layers
    w1p = np.concatenate([w1[:, 3:], w1[:, :3], np.zeros((w1.shape[0], 13), np.float32)], axis=1)
    pk = b200.ops.mlp_pack_weights
    return [(pk(dev(w1p)), dev(b1), w1.shape[0]), (pk(dev(w2)), dev(b2), w2.shape[0]), (pk(dev(w3)), dev(b3), w3.shape[0])]


def run_fused(b200, xyz, new_xyz, idx, layers, radius, rot, feat, passes=3, channels_last=False):
    if feat is None:
        return b200.ops.group_mlp_maxpool(dev(xyz), dev(new_xyz), dev(idx), device_layers(b200, layers), radius=radius,
                                          rot=None if rot is None else dev(rot), passes=passes,
                                          channels_last=channels_last)
    assert idx.shape[0] == 1 and rot is None
    feat_t = dev(np.ascontiguousarray(feat.transpose(0, 2, 1)))
    r = b200.ops.group_feat_mlp_maxpool(dev(xyz), dev(new_xyz), feat_t, dev(idx[0]), feat_layers(b200, layers),
                                        radius=radius, passes=passes, channels_last=channels_last)
    if channels_last:
        return r[0][..., None], r[1][:, :, None, :]
    return r[..., None]


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(HERE, "golden", "mlp_*.npz"))))
def test_fused_matches_reference_golden(b200, oracle_mod, path):
    z = np.load(path)
    layers = [oracle_mod.mlp.fold_bn(*p) for p in b200.synth.mlp_params(int(z["seed"]), [int(d) for d in z["dims"]])]
    rot = z["rot"] if z["rot"].size else None
    feat = z["feat"] if z["feat"].size else None
    got, got_t = run_fused(b200, z["xyz"], z["new_xyz"], z["idx"], layers, float(z["radius"]), rot, feat,
                           channels_last=True)
    got, got_t = got.cpu().numpy(), got_t.cpu().numpy()
    assert got.shape == z["out"].shape
    assert close(got, z["out"], RTOL)
    assert np.array_equal(got_t, got.transpose(0, 2, 3, 1))          # (B,M,G,C): the same values channels-last
    fast = run_fused(b200, z["xyz"], z["new_xyz"], z["idx"], layers, float(z["radius"]), rot, feat, passes=1)
    assert close(fast.cpu().numpy(), z["out"], 3e-3)


@pytest.mark.parametrize("dims,b,n,m,ns,radius", [
    ([3 + 128, 128, 128, 256], 1, 2048, 1024, 32, 0.1),          # SA2 at config 2
    ([3 + 256, 128, 128, 256], 1, 1024, 512, 16, 0.2),           # SA3
    ([3 + 256, 128, 128, 256], 2, 512, 256, 16, 0.3),            # SA4, two clouds
    ([3 + 16, 16, 32, 16], 3, 300, 7, 64, 0.5),
    ([3 + 480, 256, 128, 128], 1, 400, 33, 128, 0.4)])
def test_fused_with_features_matches_oracle(b200, oracle_mod, dims, b, n, m, ns, radius):
    seed = n + m + dims[0]
    xyz = np.stack([b200.synth.tabletop_cloud(n, seed + i) for i in range(b)])
    feat = np.random.default_rng(seed).standard_normal((b, dims[0] - 3, n)).astype(np.float32)
    xd = dev(xyz)
    inds = b200.ops.furthest_point_sampling(xd, m)
    new_xyz = b200.ops.gather_xyz(xd, inds, transposed=False)
    idx = b200.ops.ball_query(new_xyz, xd, radius, ns).cpu().numpy()[None]
    layers = [oracle_mod.mlp.fold_bn(*p) for p in b200.synth.mlp_params(seed, dims)]
    got = run_fused(b200, xyz, new_xyz.cpu().numpy(), idx, layers, radius, None, feat).cpu().numpy()
    want = oracle_mod.mlp.group_mlp_maxpool(xyz, new_xyz.cpu().numpy(), idx[0], layers, radius, None, feat)[..., None]
    assert got.shape == want.shape
    assert close(got, want, RTOL)


@pytest.mark.parametrize("dims,b,n,m,ns,g,crop", [
    ([3, 64, 64, 128], 1, 20000, 2048, 64, 1, False),          # SA1 at config 2
    ([3, 64, 128, 256], 1, 20000, 1024, 64, 4, True),          # CloudCrop at config 2
    ([3, 64, 128, 256], 2, 3000, 37, 32, 3, True),             # ragged: lists do not fill the last tile
    ([3, 16, 256, 16], 1, 500, 5, 128, 1, False),
    ([3, 256, 256, 256], 1, 2000, 50, 16, 2, True)])
def test_fused_matches_oracle(b200, oracle_mod, dims, b, n, m, ns, g, crop):
    seed = n + m
    xyz = np.stack([b200.synth.tabletop_cloud(n, seed + i) for i in range(b)])
    xd = dev(xyz)
    inds = b200.ops.furthest_point_sampling(xd, m)
    new_xyz = b200.ops.gather_xyz(xd, inds, transposed=False)
    if crop:
        rot = np.stack([b200.synth.seed_rotations(m, seed + i) for i in range(b)]).reshape(b, m, 9)
        idx = b200.ops.cylinder_query_multi(new_xyz, xd, dev(rot), 0.05, -0.02, [0.01 * (k + 1) for k in range(g)], ns)
        radius = 0.0
    else:
        rot = None
        idx = b200.ops.ball_query(new_xyz, xd, 0.04, ns)[None].contiguous()
        radius = 0.04
    layers = [oracle_mod.mlp.fold_bn(*p) for p in b200.synth.mlp_params(seed, dims)]
    got = b200.ops.group_mlp_maxpool(xd, new_xyz, idx, device_layers(b200, layers), radius=radius,
                                     rot=None if rot is None else dev(rot)).cpu().numpy()
    want = oracle_mod.mlp.group_mlp_maxpool(xyz, new_xyz.cpu().numpy(), idx.cpu().numpy(), layers, radius, rot)
    assert got.shape == want.shape
    assert close(got, want, RTOL)


def test_fused_rejects_unsupported_shapes(b200):
    x = torch.zeros(1, 64, 3, device="cuda")
    c = torch.zeros(1, 4, 3, device="cuda")
    idx = torch.zeros(1, 4, 24, dtype=torch.int32, device="cuda")                  # nsample 24
    w = torch.zeros(64, 3, device="cuda")
    bz = torch.zeros(64, device="cuda")
    pk = b200.ops.mlp_pack_weights(torch.zeros(64, 64, device="cuda"))
    with pytest.raises(RuntimeError):
        b200.ops.group_mlp_maxpool(x, c, idx, [(w, bz), (pk, bz, 64), (pk, bz, 64)])
    with pytest.raises(RuntimeError):
        b200.ops.mlp_pack_weights(torch.zeros(60, 64, device="cuda"))


def test_modules_fused_equals_composed_and_keys_match_reference(b200):
    fm = b200.fused_modules
    torch.manual_seed(0)
    crop = fm.CloudCrop(64, 3).cuda()
    sa1 = fm.PointnetSAModuleVotes(npoint=512, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], use_xyz=True,
                                   normalize_xyz=True).cuda()
    for mod in (crop, sa1):
        for m_ in mod.modules():
            if isinstance(m_, torch.nn.BatchNorm2d):
                m_.running_mean.normal_(0, 0.3)
                m_.running_var.uniform_(0.5, 2.0)
                m_.weight.data.uniform_(0.5, 1.5)
                m_.bias.data.normal_(0, 0.2)
        mod.eval()
    # the reference's parameter names (SharedMLP / _ConvBase / _BNBase: pytorch_utils.py:11-116)
    assert list(crop.state_dict())[:6] == ["mlps.layer0.conv.weight", "mlps.layer0.bn.bn.weight", "mlps.layer0.bn.bn.bias",
                                           "mlps.layer0.bn.bn.running_mean", "mlps.layer0.bn.bn.running_var",
                                           "mlps.layer0.bn.bn.num_batches_tracked"]
    assert "mlp_module.layer2.conv.weight" in sa1.state_dict()
    xyz = dev(np.stack([b200.synth.tabletop_cloud(6000, 3 + i) for i in range(2)]))
    with torch.no_grad():
        new_xyz, feat, inds = sa1(xyz)
    with torch.enable_grad():                                   # grad mode: the composed path
        new_xyz2, feat2, inds2 = sa1(xyz)
    assert torch.equal(inds, inds2) and torch.equal(new_xyz, new_xyz2)
    assert feat.shape == feat2.shape == (2, 128, 512)
    assert close(feat.cpu().numpy(), feat2.detach().cpu().numpy(), 2e-3)        # torch convs run TF32 by default
    seeds = new_xyz[:, :128].contiguous()
    rot = dev(np.stack([b200.synth.seed_rotations(128, 9 + i) for i in range(2)]).reshape(2, 128, 3, 3))
    with torch.no_grad():
        f = crop(seeds, xyz, rot)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.enable_grad():
            f2 = crop(seeds, xyz, rot)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    assert f.shape == f2.shape == (2, 256, 128, 4)
    assert close(f.cpu().numpy(), f2.detach().cpu().numpy(), 2e-5)


def test_backbone_sa_chain_fused_equals_composed(b200):
    """SA1 -> SA2 -> SA3 -> SA4 with the reference's widths (models/backbone.py:30-64), features handed from module to
    module channels-last; against the same modules' composed path with cuDNN TF32 switched off."""
    fm = b200.fused_modules
    torch.manual_seed(1)
    cfg = [(512, 0.04, 64, [0, 64, 64, 128]), (256, 0.1, 32, [128, 128, 128, 256]),
           (128, 0.2, 16, [256, 128, 128, 256]), (64, 0.3, 16, [256, 128, 128, 256])]
    mods = [fm.PointnetSAModuleVotes(npoint=n, radius=r, nsample=ns, mlp=list(mlp), use_xyz=True, normalize_xyz=True).cuda()
            for n, r, ns, mlp in cfg]
    for mod in mods:
        for m_ in mod.modules():
            if isinstance(m_, torch.nn.BatchNorm2d):
                m_.running_mean.normal_(0, 0.3)
                m_.running_var.uniform_(0.5, 2.0)
                m_.weight.data.uniform_(0.5, 1.5)
                m_.bias.data.normal_(0, 0.2)
        mod.eval()
    xyz0 = dev(np.stack([b200.synth.tabletop_cloud(5000, 11 + i) for i in range(2)]))
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        xyz, feat, feat_t = xyz0, None, None
        cxyz, cfeat = xyz0, None
        for mod in mods:
            with torch.no_grad():
                xyz, feat, inds = mod(xyz, feat, features_t=feat_t)
            feat_t = mod.features_t
            assert feat_t is not None and torch.equal(feat_t, feat.transpose(1, 2))
            with torch.enable_grad():
                cxyz, cfeat, cinds = mod(cxyz, cfeat)
            assert torch.equal(inds, cinds) and torch.equal(xyz, cxyz)
            assert close(feat.cpu().numpy(), cfeat.detach().cpu().numpy(), 5e-5)
            cfeat = cfeat.detach()
    finally:
        torch.backends.cudnn.allow_tf32 = old
    assert feat.shape == (2, 256, 64)
