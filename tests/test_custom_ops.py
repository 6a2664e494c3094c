"""torch.ops.b200grasp.* (graspnet-baseline_b200/custom_ops.py): the C-ABI calls registered as torch custom operators.
CPU: schemas and FAKE implementations (shape / dtype propagation under FakeTensorMode, no kernel runs).
GPU: torch.library.opcheck (schema, fake tensor, autograd registration, AOT dispatch) and equality with the
drop-in functions of pointnet2_utils."""
import pytest
import torch


def fake_inputs(device):
    xyz = torch.empty(2, 500, 3, device=device)
    new_xyz = torch.empty(2, 64, 3, device=device)
    feats = torch.empty(2, 16, 500, device=device)
    idx = torch.empty(2, 64, dtype=torch.int32, device=device)
    gidx = torch.empty(2, 64, 8, dtype=torch.int32, device=device)
    rot = torch.empty(2, 64, 9, device=device)
    nn_idx = torch.empty(2, 64, 3, dtype=torch.int32, device=device)
    w = torch.empty(2, 64, 3, device=device)
    return xyz, new_xyz, feats, idx, gidx, rot, nn_idx, w


def test_schemas_and_fake_implementations(b200):
    from torch._subclasses.fake_tensor import FakeTensorMode
    o = torch.ops.b200grasp
    for name in b200.custom_ops.OPS:
        assert hasattr(o, name), name
    assert "Tensor points, SymInt nsamples) -> Tensor" in str(o.furthest_point_sampling.default._schema)
    with FakeTensorMode():
        xyz, new_xyz, feats, idx, gidx, rot, nn_idx, w = fake_inputs("cuda")
        r = o.furthest_point_sampling(xyz, 64)
        assert r.shape == (2, 64) and r.dtype == torch.int32 and r.device.type == "cuda"
        assert o.gather_points(feats, idx).shape == (2, 16, 64)
        assert o.gather_points_grad(torch.empty(2, 16, 64, device="cuda"), idx, 500).shape == (2, 16, 500)
        assert o.ball_query(new_xyz, xyz, 0.1, 8).shape == (2, 64, 8)
        assert o.cylinder_query(new_xyz, xyz, rot, 0.05, -0.02, 0.04, 8).dtype == torch.int32
        assert o.group_points(feats, gidx).shape == (2, 16, 64, 8)
        assert o.group_points_grad(torch.empty(2, 16, 64, 8, device="cuda"), gidx, 500).shape == (2, 16, 500)
        d, i = o.three_nn(new_xyz, xyz)
        assert d.shape == (2, 64, 3) and i.dtype == torch.int32
        assert o.three_interpolate(feats, nn_idx, w).shape == (2, 16, 64)
        assert o.three_interpolate_grad(torch.empty(2, 16, 64, device="cuda"), nn_idx, w, 500).shape == (2, 16, 500)
        k = o.knn(torch.empty(1, 3, 300, device="cuda"), torch.empty(1, 3, 40, device="cuda"), 1)
        assert k.shape == (1, 1, 40) and k.dtype == torch.int64
        # autograd is registered: outputs of the differentiable ops require grad under fake tensors too
        f = torch.empty(2, 16, 500, device="cuda", requires_grad=True)
        assert o.gather_points(f, idx).requires_grad and o.group_points(f, gidx).requires_grad
        assert o.three_interpolate(f, nn_idx, w).requires_grad


@pytest.mark.gpu
def test_opcheck_and_equality_with_dropins(b200):
    from torch.library import opcheck
    from b200grasp import pointnet2_utils as pu
    o = torch.ops.b200grasp
    g = torch.Generator(device="cuda").manual_seed(0)
    xyz = torch.rand(2, 500, 3, device="cuda", generator=g)
    feats = torch.randn(2, 16, 500, device="cuda", generator=g, requires_grad=True)
    inds = o.furthest_point_sampling(xyz, 64)
    assert torch.equal(inds, pu.furthest_point_sample(xyz, 64))
    new_xyz = pu.gather_operation(xyz.transpose(1, 2).contiguous(), inds).transpose(1, 2).contiguous()
    gidx = o.ball_query(new_xyz, xyz, 0.2, 8)
    assert torch.equal(gidx, pu.ball_query(0.2, 8, xyz, new_xyz))
    dist2, nn_idx = o.three_nn(new_xyz, xyz)
    w = torch.softmax(-dist2, dim=2)
    checks = [(o.furthest_point_sampling, (xyz, 64)), (o.ball_query, (new_xyz, xyz, 0.2, 8)),
              (o.gather_points, (feats, inds)), (o.group_points, (feats, gidx)), (o.three_nn, (new_xyz, xyz)),
              (o.three_interpolate, (feats, nn_idx, w)),
              (o.knn, (torch.rand(1, 3, 300, device="cuda"), torch.rand(1, 3, 40, device="cuda"), 1))]
    for op, args in checks:
        opcheck(op, args, test_utils=("test_schema", "test_faketensor", "test_autograd_registration",
                                      "test_aot_dispatch_static", "test_aot_dispatch_dynamic"))
    # gradients through the custom ops == through the reference-structured autograd.Functions
    for fn_op, fn_ref in ((lambda f: o.gather_points(f, inds), lambda f: pu.gather_operation(f, inds)),
                          (lambda f: o.group_points(f, gidx), lambda f: pu.grouping_operation(f, gidx)),
                          (lambda f: o.three_interpolate(f, nn_idx, w), lambda f: pu.three_interpolate(f, nn_idx, w))):
        a = feats.detach().clone().requires_grad_(True)
        b = feats.detach().clone().requires_grad_(True)
        ya, yb = fn_op(a), fn_ref(b)
        assert torch.equal(ya, yb)
        cot = torch.randn_like(ya)
        ya.backward(cot)
        yb.backward(cot)
        assert (a.grad - b.grad).abs().max() <= 1e-5 * b.grad.abs().max()
    # torch.compile sees through the operator (no graph break on it)
    compiled = torch.compile(lambda f, i: torch.ops.b200grasp.gather_points(f, i) * 2.0, fullgraph=True)
    assert torch.equal(compiled(feats.detach(), inds), o.gather_points(feats.detach(), inds) * 2.0)
