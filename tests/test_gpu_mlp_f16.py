"""fp16x3 tensor-core kernels (csrc/mlp_f16.cu) through the C ABI: stacks of 1x1 convolutions over point columns
(reference: PointnetFPModule's SharedMLP pointnet2_modules.py:409-414, ApproachNet models/modules.py:52-57, the heads
:170-177 / :215-222) against an fp64 statement of the same layers, and the grouped kernels against the 3xTF32 kernel.
Tolerance: 1e-5 of the tensor's scale (max |err| / max |want|), the bound the module-level golden tests use."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
REL = 1e-5


def rel(got, want):
    return float((got.double() - want).abs().max() / want.abs().max())


def make_layers(dims, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return [(torch.randn(co, ci, device="cuda", generator=g) * (2.0 / ci) ** 0.5 * scale,
             torch.randn(co, device="cuda", generator=g) * 0.1) for ci, co in zip(dims[:-1], dims[1:])]


def reference(x, ws, relu_last):
    y = x.double()
    for i, (w, b) in enumerate(ws):
        y = torch.einsum("oc,bcn->bon", w.double(), y) + b.double()[None, :, None]
        if relu_last or i + 1 < len(ws):
            y = torch.relu(y)
    return y


@pytest.mark.parametrize("dims,b,n,c2,relu_last", [
    ([512, 256, 256], 2, 1024, 256, True),          # FP module: [interpolated | skip] -> 256 -> 256
    ([256, 256, 302, 302], 1, 1024, 0, False),      # ApproachNet: a 302-wide layer split over two heads of a launch
    ([256, 128, 128, 36], 1, 4096, 0, False),       # OperationNet
    ([256, 128, 128, 12], 3, 1000, 0, False),       # ToleranceNet, ragged columns (n % 128 != 0)
    ([256, 302], 64, 1024, 0, False),               # single wide layer, 3-4 tiles per CTA, source longer than the ring
    ([40, 16], 1, 130, 0, True),                    # narrow, K padded 40 -> 64, two tiles
])
def test_pointwise_stack_f16_matches_fp64(b200, dims, b, n, c2, relu_last):
    from b200grasp import graspnet, fused_modules as fm
    ws = make_layers(dims, 7)
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.relu(torch.randn(b, dims[0], n, device="cuda", generator=g)) * 2
    x1, x2 = (x[:, :dims[0] - c2].contiguous(), x[:, dims[0] - c2:].contiguous()) if c2 else (x, None)
    fold = graspnet._Folded(torch.nn.Module(), lambda: ws, relu_last=relu_last)
    assert fm.ARITH == "f16x3"
    got = graspnet._stack(x1, fold, x2=x2)
    torch.cuda.synchronize()
    want = reference(x, ws, relu_last)
    assert got.shape == want.shape
    assert rel(got, want) <= REL


def test_pointwise_stack_f16_range_and_special_values(b200):
    """the power-of-two scaling keeps fp16 operands in range whatever the magnitudes: huge and tiny inputs / weights;
    NaN stays in its column"""
    from b200grasp import graspnet
    for xs, wsn in ((1e4, 1.0), (1e-6, 1.0), (1.0, 300.0), (3e3, 1e-3)):
        ws = make_layers([64, 48, 32], 11, scale=wsn)
        g = torch.Generator(device="cuda").manual_seed(5)
        x = torch.randn(1, 64, 256, device="cuda", generator=g) * xs
        fold = graspnet._Folded(torch.nn.Module(), lambda: ws, relu_last=False)
        got = graspnet._stack(x, fold)
        want = reference(x, ws, False)
        assert torch.isfinite(got).all()
        assert rel(got, want) <= REL, (xs, wsn)
    ws = make_layers([32, 32], 2)
    x = torch.randn(1, 32, 256, device="cuda")
    x[0, 5, 77] = float("nan")
    got = graspnet._stack(x, graspnet._Folded(torch.nn.Module(), lambda: ws, relu_last=False))
    bad = torch.isnan(got).any(dim=1)[0]
    assert bad[77] and torch.isfinite(got[0, :, 128:]).all()          # the NaN column; other warps' rows untouched


def test_two_heads_one_launch_equals_separate_launches(b200):
    from b200grasp import ops
    wa, wt = make_layers([256, 128, 128, 36], 21), make_layers([256, 128, 128, 12], 22)
    x = torch.relu(torch.randn(2, 256, 4096, device="cuda"))
    pads = [128, 128, 48]
    la, lt = ops.f16_stack(wa, relu_last=False, n_pads=pads), ops.f16_stack(wt, relu_last=False, n_pads=pads)
    ya, yt = ops.pointwise_stack_f16(x, la, layers2=lt)
    sa = ops.pointwise_stack_f16(x, ops.f16_stack(wa, relu_last=False))
    st = ops.pointwise_stack_f16(x, ops.f16_stack(wt, relu_last=False))
    assert torch.equal(ya, sa) and torch.equal(yt, st)
    assert rel(ya, reference(x, wa, False)) <= REL and rel(yt, reference(x, wt, False)) <= REL


@pytest.mark.parametrize("batch", [1, 5])
def test_grouped_f16_agrees_with_tf32x3_kernel(b200, batch):
    """SA1 / CloudCrop (xyz_pipe_kernel) and SA2 (features) in fp16x3 against the round-1 3xTF32 kernel: both are
    fp32-grade, so they agree to 1e-5 of scale; several tiles per CTA at batch 5"""
    from b200grasp import ops, synth, fused_modules as fm
    xyz = torch.from_numpy(np.stack([synth.tabletop_cloud(20000, 5 + i) for i in range(batch)])).cuda()
    inds = ops.furthest_point_sampling(xyz, 2048)
    new_xyz = ops.gather_xyz(xyz, inds, transposed=False)
    with torch.no_grad():
        crop = fm.CloudCrop(64, 3).cuda().eval()
        synth.fill_state_dict(crop, 3)
        seeds = new_xyz[:, :1024].contiguous()
        rot = torch.from_numpy(np.stack([synth.seed_rotations(1024, 9 + i) for i in range(batch)]).reshape(batch, 1024, 9)).cuda()
        idx = ops.cylinder_query_multi(seeds, xyz, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64)
        first, l16 = crop._fused.get_f16()
        a = ops.group_mlp_maxpool_f16(xyz, seeds, idx, first, l16, rot=rot)
        b = ops.group_mlp_maxpool(xyz, seeds, idx, crop._fused.get(), rot=rot, passes=3)
        assert rel(a, b.double()) <= REL
        sa2 = fm.PointnetSAModuleVotes(npoint=1024, radius=0.1, nsample=32, mlp=[128, 128, 128, 256], normalize_xyz=True).cuda().eval()
        synth.fill_state_dict(sa2, 5)
        idx2 = ops.ball_query(seeds, new_xyz, 0.1, 32)
        feat_t = torch.relu(torch.randn(batch, 2048, 128, device="cuda")) * 3
        a = ops.group_feat_mlp_maxpool_f16(new_xyz, seeds, feat_t, idx2, sa2._fused.get_f16()[1], radius=0.1)
        b = ops.group_feat_mlp_maxpool(new_xyz, seeds, feat_t, idx2, sa2._fused.get(), radius=0.1, passes=3)
        assert rel(a, b.double()) <= REL


@pytest.mark.parametrize("batch", [1, 3])
def test_row_bound_rider_gives_the_same_bits_as_the_row_scan(b200, batch):
    """the grouped fp16x3 kernels leave each output row's largest magnitude beside the channels-last output
    (out_t.b200_row_bound); the next module reads that instead of scanning every gathered row.  SA1 -> SA2 -> SA3 with
    the rider against the same chain on detached copies of the features (no rider): equal bits; the rider itself
    equals the row maxima"""
    from b200grasp import ops, synth, fused_modules as fm
    xyz = torch.from_numpy(np.stack([synth.tabletop_cloud(20000, 50 + i) for i in range(batch)])).cuda()
    inds = ops.furthest_point_sampling(xyz, 2048)
    p1 = ops.gather_xyz(xyz, inds, transposed=False)
    with torch.no_grad():
        sa1 = fm.PointnetSAModuleVotes(npoint=2048, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], normalize_xyz=True).cuda().eval()
        sa2 = fm.PointnetSAModuleVotes(npoint=1024, radius=0.1, nsample=32, mlp=[128, 128, 128, 256], normalize_xyz=True).cuda().eval()
        sa3 = fm.PointnetSAModuleVotes(npoint=512, radius=0.2, nsample=16, mlp=[256, 128, 128, 256], normalize_xyz=True).cuda().eval()
        for i, sa in enumerate((sa1, sa2, sa3)):
            synth.fill_state_dict(sa, 70 + i)
        p2, p3 = p1[:, :1024].contiguous(), p1[:, :512].contiguous()
        i1 = ops.ball_query(p1, xyz, 0.04, 64)
        i2 = ops.ball_query(p2, p1, 0.1, 32)
        i3 = ops.ball_query(p3, p2, 0.2, 16)
        f1, t1 = sa1.pooled(xyz, p1, i1)
        rb1 = t1.b200_row_bound
        assert rb1.shape == (batch * 2048, 8)
        assert torch.equal(rb1.max(dim=1).values.view(batch, 2048), t1.abs().amax(dim=2))
        f2, t2 = sa2.pooled(p1, p2, i2, features_t=t1)
        assert torch.equal(t2.b200_row_bound.max(dim=1).values.view(batch, 1024), t2.abs().amax(dim=2))
        f3, t3 = sa3.pooled(p2, p3, i3, features_t=t2)
        g2, u2 = sa2.pooled(p1, p2, i2, features_t=t1.clone())          # a copy carries no rider: the row scan runs
        g3, u3 = sa3.pooled(p2, p3, i3, features_t=t2.clone())
        assert not hasattr(t1.clone(), "b200_row_bound")
        assert torch.equal(f2, g2) and torch.equal(t2, u2) and torch.equal(f3, g3) and torch.equal(t3, u3)


def test_tiles_per_cta_knob_does_not_change_results(b200):
    """b200_mlp_f16_set_tiles_per_cta (GraspNetPipeline runs its slots' few-tile launches with 4 tiles per CTA): SA3-like
    grouped module (64 tiles) and two heads (32 tiles each) with 1, 2, 4 and 64 tiles per CTA give the same bits"""
    from b200grasp import ops, synth, _abi, fused_modules as fm
    lib = _abi.lib
    xyz = torch.from_numpy(synth.tabletop_cloud(20000, 3)[None]).cuda()
    pts = ops.gather_xyz(xyz, ops.furthest_point_sampling(xyz, 1024), transposed=False)
    wa, wt = make_layers([256, 128, 128, 36], 21), make_layers([256, 128, 128, 12], 22)
    x = torch.relu(torch.randn(1, 256, 4096, device="cuda"))
    la, lt = ops.f16_stack(wa, relu_last=False, n_pads=[128, 128, 48]), ops.f16_stack(wt, relu_last=False, n_pads=[128, 128, 48])
    with torch.no_grad():
        sa = fm.PointnetSAModuleVotes(npoint=512, radius=0.2, nsample=16, mlp=[256, 128, 128, 256], normalize_xyz=True).cuda().eval()
        synth.fill_state_dict(sa, 9)
        seeds = pts[:, :512].contiguous()
        idx = ops.ball_query(seeds, pts, 0.2, 16)
        feat_t = torch.relu(torch.randn(1, 1024, 256, device="cuda")) * 3
        l16 = sa._fused.get_f16()[1]
        sa1 = fm.PointnetSAModuleVotes(npoint=1024, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], normalize_xyz=True).cuda().eval()
        synth.fill_state_dict(sa1, 10)
        idx1 = ops.ball_query(pts, xyz, 0.04, 64)
        first1, l1 = sa1._fused.get_f16()
        res = []
        try:
            for k in (1, 2, 4, 64):
                lib.b200_mlp_f16_set_tiles_per_cta(k)
                a = ops.group_feat_mlp_maxpool_f16(pts, seeds, feat_t, idx, l16, radius=0.2)
                ya, yt = ops.pointwise_stack_f16(x, la, layers2=lt)
                ya = torch.cat([ya.flatten(), ops.group_mlp_maxpool_f16(xyz, pts, idx1, first1, l1, radius=0.04).flatten()])
                torch.cuda.synchronize()
                res.append((a.clone(), ya.clone(), yt.clone()))
        finally:
            lib.b200_mlp_f16_set_tiles_per_cta(1)
    for r in res[1:]:
        assert all(torch.equal(u, v) for u, v in zip(res[0], r))
