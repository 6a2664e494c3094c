"""The whole GraspNet forward (graspnet-baseline_b200/graspnet.py) against golden outputs of the REFERENCE network
(tests/golden/graspnet_*.npz: models/graspnet.py::GraspNet + pred_decode run unmodified on the CPU, pointnet2 ops
answered by the C oracle -- tests/golden/make_golden.py), the fused path against the composed path, and the
CUDA-graph engine against the call-by-call sequence.

Tolerances: sample / seed indices bit-exact; features 1e-5 relative to the tensor's scale PER MODULE (north_star:
"features within 1e-5 relative (fp32)"; every fused module alone is held to 1e-5 in tests/test_gpu_mlp_fused.py).
Through the network the bound is 1e-5 x the number of fused grouped-MLP modules chained in front of the tensor
(SA4 / FP2 / ApproachNet outputs: 4; CloudCrop and the grasp heads: 1): the tensor core adds each K = 8 product group
to its fp32 accumulator with truncation, a bias of up to half an ulp per step that a chain of modules accumulates (the
measured errors are written to gpurun_out/graspnet_parity.txt; the reference's own GPU arithmetic -- cuDNN
convolutions with TF32 allowed, torch's default -- is ~1e-3 from the same fp32 CPU results).  Discrete selections
downstream of features (best view, decoded classes) must agree on all but a fraction of seeds (near-ties flip).
"""
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(__file__), "golden")
REL = 1e-5


def _net(b200, seed, **kw):
    net = b200.graspnet.GraspNet(input_feature_dim=0, num_view=300, num_angle=12, num_depth=4, cylinder_radius=0.05,
                                 hmin=-0.02, hmax_list=[0.01, 0.02, 0.03, 0.04], is_training=False, **kw)
    b200.synth.fill_state_dict(net, seed)
    return net


def close(got, want, rel=REL, what=""):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    scale = np.abs(want).max() if want.size else 1.0
    err = np.abs(got - want).max() / max(scale, 1e-30) if want.size else 0.0
    out = os.path.join(os.path.dirname(os.path.dirname(__file__)), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "graspnet_parity.txt"), "a") as f:
            f.write("%-40s max error %.3g of scale (tolerance %.1g)\n" % (what, err, rel))
    assert err <= rel, "%s: max error %.3g of scale (tolerance %.1g)" % (what, err, rel)


@pytest.mark.parametrize("name", ["a", "b"])
def test_state_dict_keys_are_the_reference_s(name):
    """CPU: same parameter / buffer names and shapes as the reference GraspNet -> its checkpoints load unchanged"""
    import b200grasp
    g = np.load(os.path.join(GOLD, "graspnet_%s.npz" % name))
    sd = b200grasp.graspnet.GraspNet(is_training=False).state_dict()
    assert sorted(sd) == list(g["state_keys"])
    assert [",".join(map(str, sd[k].shape)) for k in sorted(sd)] == list(g["state_shapes"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["a", "b"])
def test_forward_matches_reference_network(b200, name):
    g = np.load(os.path.join(GOLD, "graspnet_%s.npz" % name))
    net = _net(b200, int(g["seed"])).cuda().eval()
    with torch.no_grad():
        ep = net({"point_clouds": torch.from_numpy(g["cloud"]).cuda()})
        rows = b200.graspnet.pred_decode(ep)
    cpu = lambda k: ep[k].cpu().numpy()
    assert (cpu("sa1_inds") == g["sa1_inds"]).all() and (cpu("fp2_inds") == g["fp2_inds"]).all()
    close(cpu("sa1_features")[:, :, ::16], g["sa1_features"], what="golden %s sa1_features" % name)
    loose = 1.0 if b200.fused_modules.STACK_ARITH == "fp32" else 4.0
    close(cpu("sa4_features")[:, :, ::4], g["sa4_features"], REL, what="sa4_features")
    close(cpu("fp2_features")[:, :, ::8], g["fp2_features"], loose * REL, what="fp2_features")
    close(cpu("objectness_score"), g["objectness_score"], max(loose, 1.3) * REL, what="objectness_score")
    close(cpu("view_score")[:, ::8], g["view_score"], loose * REL, what="view_score")
    same_view = cpu("grasp_top_view_inds") == g["grasp_top_view_inds"]
    assert same_view.mean() >= 0.995, "best view differs on %d seeds" % (~same_view).sum()
    close(cpu("grasp_top_view_score"), g["grasp_top_view_score"], loose * REL, what="top view score")
    assert np.abs(cpu("grasp_top_view_xyz") - g["grasp_top_view_xyz"])[same_view].max() <= 1e-6
    assert np.abs(cpu("grasp_top_view_rot") - g["grasp_top_view_rot"])[same_view].max() <= 1e-6
    sub = same_view[:, ::4]
    for k in ("grasp_score_pred", "grasp_angle_cls_pred", "grasp_width_pred", "grasp_tolerance_pred"):
        got, want = cpu(k)[:, :, ::4], g[k]
        sel = np.broadcast_to(sub[:, None, :, None], want.shape)
        close(got[sel], want[sel], what=k)
    for i, r in enumerate(rows):
        want = g["rows_%d" % i]
        got = r.cpu().numpy()
        assert got.shape == want.shape, "scene %d: %d rows decoded, reference %d" % (i, got.shape[0], want.shape[0])
        # rows are in seed order; classes (depth, angle) are argmaxes and may flip on near-ties -> row-wise agreement
        ok = np.abs(got - want).max(axis=1) <= 1e-4
        assert ok.mean() >= 0.98, "scene %d: %d of %d decoded rows differ" % (i, (~ok).sum(), len(ok))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["a", "b"])
def test_error_against_exact_network(b200, name):
    """How far is this package's float32 forward from the EXACT network (the reference evaluated in float64 on the
    same indices, golden key 'exact:*'), next to the reference's own float32 CPU run?  Metric everywhere in this file:
    max |err| over the tensor / max |exact| over the tensor ("of scale"), not per-element relative error.
    The reference's fp32 run sits at 0.6-2.8e-6; the tensor-core path (fp16x3: three fp16 products per K = 16 step,
    the accumulator updated with truncation 3 K / 16 times per output) at 0.2-2e-5, largest on the two objectness
    channels at the end of the SA1..SA4 -> FP -> ApproachNet chain."""
    g = np.load(os.path.join(GOLD, "graspnet_%s.npz" % name))
    net = _net(b200, int(g["seed"])).cuda().eval()
    with torch.no_grad():
        ep = net({"point_clouds": torch.from_numpy(g["cloud"]).cuda()})
    cpu = lambda k: ep[k].cpu().numpy()
    out = os.path.join(os.path.dirname(os.path.dirname(__file__)), "gpurun_out")
    if b200.fused_modules.STACK_ARITH == "fp32":        # the default: FP / ApproachNet stacks in exact fp32
        bound = {"sa4_features": 1e-5, "fp2_features": 1e-5, "objectness_score": 1.2e-5, "view_score": 1e-5,
                 "grasp_top_view_score": 1e-5}
    else:                                               # all stacks on the tensor cores (fp16x3)
        bound = {"sa4_features": 1e-5, "fp2_features": 1.5e-5, "objectness_score": 3e-5, "view_score": 2e-5,
                 "grasp_top_view_score": 2e-5}
    sl = {"sa4_features": lambda x: x[:, :, ::4], "fp2_features": lambda x: x[:, :, ::8], "objectness_score": lambda x: x,
          "view_score": lambda x: x[:, ::8], "grasp_top_view_score": lambda x: x}
    for k in bound:
        exact = g["exact:" + k].astype(np.float64)
        scale = np.abs(exact).max()
        ours = np.abs(sl[k](cpu(k)).astype(np.float64) - exact).max() / scale
        ref = np.abs(g[k].astype(np.float64) - exact).max() / scale
        if os.path.isdir(out):
            with open(os.path.join(out, "graspnet_parity.txt"), "a") as f:
                f.write("exact %s %-28s this package %.3g   reference fp32 (CPU) %.3g   of scale\n" % (name, k, ours, ref))
        assert ours <= bound[k], "%s: %.3g of scale from the exact network (bound %.1g)" % (k, ours, bound[k])
    same = (cpu("grasp_top_view_inds") == g["exact:grasp_top_view_inds"]).mean()
    assert same >= 0.995


@pytest.mark.gpu
def test_fused_path_equals_composed_path(b200):
    """fast path (fused grouped MLPs, folded BatchNorm, view table) == the reference's op sequence on the drop-ins"""
    net = _net(b200, 90).cuda().eval()
    cloud = torch.from_numpy(np.stack([b200.synth.tabletop_cloud(6000, 5), b200.synth.tabletop_cloud(6000, 6)])).cuda()
    tf32 = torch.backends.cudnn.allow_tf32
    with torch.no_grad():
        fast = net({"point_clouds": cloud})
        b200.fused_modules.FUSED = False
        torch.backends.cudnn.allow_tf32 = False                 # the composed path's convolutions in fp32
        try:
            slow = net({"point_clouds": cloud})
        finally:
            b200.fused_modules.FUSED = True
            torch.backends.cudnn.allow_tf32 = tf32
    assert torch.equal(fast["sa1_inds"], slow["sa1_inds"])
    for depth, k in ((1, "sa1_features"), (2, "sa2_features"), (3, "sa3_features"), (4, "sa4_features"),
                     (4, "fp2_features"), (4, "objectness_score")):
        close(fast[k].cpu().numpy(), slow[k].cpu().numpy(), depth * REL, what="fused vs composed " + k)
    same = (fast["grasp_top_view_inds"] == slow["grasp_top_view_inds"])
    assert same.float().mean() >= 0.995
    d = (fast["grasp_top_view_rot"] - slow["grasp_top_view_rot"]).abs()[same]
    assert d.max() <= 1e-6
    sel = same[:, None, :, None].expand_as(fast["grasp_score_pred"])
    for k in ("grasp_score_pred", "grasp_angle_cls_pred", "grasp_width_pred", "grasp_tolerance_pred"):
        close(fast[k][sel].cpu().numpy(), slow[k][sel].cpu().numpy(), what=k)


@pytest.mark.gpu
@pytest.mark.parametrize("batch", [1, 3])
def test_engine_graph_equals_call_sequence(b200, batch):
    """GraspNetEngine (one graph, several streams, device-resident post-processing) == net() + pred_decode +
    ModelFreeCollisionDetector.detect + GraspGroup.nms called one after the other (test.py:86-96)"""
    net = _net(b200, 72).cuda().eval()
    n = 8000
    clouds = np.stack([b200.synth.tabletop_cloud(n, 20 + s) for s in range(batch)])
    eng = b200.graspnet.GraspNetEngine(net, batch=batch, num_point=n)
    res = eng.infer(clouds)
    res2 = eng.infer(clouds)                                    # replay
    with torch.no_grad():
        ep = net({"point_clouds": torch.from_numpy(clouds).cuda()})
        rows = b200.graspnet.pred_decode(ep)
    P = b200.graspnet.POST
    for s in range(batch):
        assert torch.equal(eng.end_points["sa1_inds"], ep["sa1_inds"])
        assert torch.equal(eng.end_points["fp2_features"], ep["fp2_features"]), "streams changed the arithmetic"
        gg = b200.GraspGroup(rows[s].cpu().numpy())
        det = b200.ModelFreeCollisionDetector(clouds[s], voxel_size=P["voxel_size"])
        mask = det.detect(gg, approach_dist=P["approach_dist"], collision_thresh=P["collision_thresh"])
        want = gg[~mask].nms().grasp_group_array
        assert res[s]["nrows"] == len(gg)
        assert (res[s]["mask"] == mask).all()
        assert res[s]["kept"].shape == want.shape and (res[s]["kept"] == want).all()
        assert (res2[s]["kept"] == res[s]["kept"]).all()
    # a new cloud through the same graph
    clouds2 = np.stack([b200.synth.tabletop_cloud(n, 40 + s) for s in range(batch)])
    r3 = eng.infer(clouds2)
    with torch.no_grad():
        ep2 = net({"point_clouds": torch.from_numpy(clouds2).cuda()})
        rows2 = b200.graspnet.pred_decode(ep2)
    for s in range(batch):
        assert r3[s]["nrows"] == rows2[s].shape[0]
    assert eng.launches > 0


@pytest.mark.gpu
@pytest.mark.parametrize("b,c1,c2,n,cout,relu", [(1, 256, 256, 512, 256, True), (2, 256, 0, 1024, 302, True),
                                                  (1, 302, 0, 1024, 302, False), (3, 128, 0, 4096, 36, False),
                                                  (1, 7, 5, 33, 3, True), (2, 16, 0, 5, 70, False)])
def test_pointwise_conv_matches_fp64_statement(b200, b, c1, c2, n, cout, relu):
    """csrc/pointwise.cu: y = [relu](w [x1 | x2] + bias) against the same product in fp64 (1e-6 of the scale)"""
    g = torch.Generator().manual_seed(b * 1000 + cout)
    x1 = torch.randn(b, c1, n, generator=g)
    x2 = torch.randn(b, c2, n, generator=g) if c2 else None
    w = torch.randn(cout, c1 + c2, generator=g) / (c1 + c2) ** 0.5
    bias = torch.randn(cout, generator=g)
    x = torch.cat([x1, x2], 1) if c2 else x1
    want = torch.matmul(w.double(), x.double()) + bias.double()[None, :, None]
    if relu:
        want = want.clamp_min(0)
    got = b200.ops.pointwise_conv(x1.cuda(), w.cuda(), bias.cuda(), relu=relu, x2=None if x2 is None else x2.cuda())
    assert got.shape == (b, cout, n)
    close(got.cpu().numpy(), want.numpy(), 1e-6, what="pointwise_conv %s" % ((b, c1, c2, n, cout),))


@pytest.mark.gpu
@pytest.mark.parametrize("S,V", [(1000, 300), (1003, 300), (93, 120)])
def test_view_select_is_torch_max_plus_the_reference_frames(b200, S, V):
    """first maximum over the view scores (ties, NaN) + table lookup == torch.max + the reference construction; seed
    counts that do not fill the last CTA's 8 seeds, fewer views than 32 slices x 4"""
    g = torch.Generator().manual_seed(5)
    B = 2
    f = torch.randn(B, 2 + V, S, generator=g)
    f[:, 2:, ::7] = torch.round(f[:, 2:, ::7])              # exact ties between views
    f[0, 40, 3] = float("nan")
    f[0, 90, 3] = float("nan")
    f[1, 2:, 11] = 0.5                                       # every view equal
    vp = b200.graspnet.ApproachNet(V, 256)
    views, frames = vp.tables(torch.device("cuda"))
    inds, score, xyz, rot = b200.ops.view_select(f.cuda(), views, frames)
    want_score, want_inds = torch.max(f[:, 2:].transpose(1, 2).contiguous().cuda(), dim=2)
    nan = torch.isnan(want_score)
    assert torch.equal(inds[~nan], want_inds[~nan]) and torch.equal(score[~nan], want_score[~nan])
    assert int(inds[0, 3]) == 38 and torch.isnan(score[0, 3])           # the FIRST NaN (view 40 - 2)
    assert int(inds[1, 11]) == 0
    want_xyz = views[inds]
    angle = torch.zeros(B * S, device="cuda")
    want_rot = b200.graspnet.view_frames(-want_xyz.view(-1, 3), angle).view(B, S, 3, 3)
    assert torch.equal(xyz, want_xyz)
    assert (rot - want_rot).abs().max() <= 1e-6


@pytest.mark.gpu
def test_pipeline_of_scenes_equals_one_scene_at_a_time(b200):
    """GraspNetPipeline (several scenes in flight on their own streams / graphs) returns, scene by scene, exactly
    what a single engine returns for that scene alone"""
    net = _net(b200, 72).cuda().eval()
    n = 6000
    clouds = [b200.synth.tabletop_cloud(n, 60 + s) for s in range(7)]
    eng = b200.graspnet.GraspNetEngine(net, batch=1, num_point=n)
    want = [eng.infer(c[None])[0] for c in clouds]
    pipe = b200.graspnet.GraspNetPipeline(net, depth=3, num_point=n)
    for _ in range(2):
        got = pipe.run(clouds)
        assert len(got) == len(want)
        for g, w in zip(got, want):
            assert g["nrows"] == w["nrows"] and (g["mask"] == w["mask"]).all()
            assert g["kept"].shape == w["kept"].shape and (g["kept"] == w["kept"]).all()


@pytest.mark.gpu
def test_backbone_training_step_gradients_match_legacy_ops(b200, legacy):
    """cfg5's per-replica step on the backbone: training mode (batch statistics, autograd through grouping_operation /
    three_interpolate backward) with this library's ops vs the SAME modules on the reference's own CUDA extension
    (oracle/_ref): loss equal, every parameter gradient within 1e-4 of its scale (both sides accumulate with atomics)."""
    if legacy is None:
        pytest.skip("oracle/_ref (legacy extension) not built")
    from b200grasp import pointnet2_utils
    cloud = torch.from_numpy(np.stack([b200.synth.tabletop_cloud(6000, 30 + s) for s in range(2)])).cuda()
    target = torch.randn(2, 256, 1024, generator=torch.Generator().manual_seed(1)).cuda()
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    grads, losses = [], []
    try:
        for ext in (None, legacy["ext"]):
            bb = b200.graspnet.Pointnet2Backbone()
            b200.synth.fill_state_dict(bb, 7)
            bb = bb.cuda().train()
            saved = pointnet2_utils._ext
            if ext is not None:
                pointnet2_utils._ext = ext
            try:
                feats, _, _ = bb(cloud)
                loss = ((feats - target) ** 2).mean()
                loss.backward()
            finally:
                pointnet2_utils._ext = saved
            losses.append(loss.item())
            grads.append({k: p.grad.detach().clone() for k, p in bb.named_parameters()})
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
    assert abs(losses[0] - losses[1]) <= 1e-5 * abs(losses[1])
    assert all(g.abs().max() > 0 for k, g in grads[0].items() if "sa" in k and "weight" in k and "bn" not in k)
    for k in grads[0]:
        close(grads[0][k].cpu().numpy(), grads[1][k].cpu().numpy(), 1e-4, what="grad " + k)


@pytest.mark.gpu
def test_pipeline_refuses_to_overwrite_an_uncollected_result(b200):
    net = _net(b200, 12).cuda().eval()
    pipe = b200.graspnet.GraspNetPipeline(net, depth=2, num_point=4000)
    cloud = b200.synth.tabletop_cloud(4000, 3)
    a = pipe.submit(cloud)
    b_ = pipe.submit(cloud)
    with pytest.raises(RuntimeError):
        pipe.submit(cloud)                                   # slot `a` again: its host result was never collected
    ra, rb = pipe.collect(a), pipe.collect(b_)
    assert ra["nrows"] == rb["nrows"] and (ra["kept"] == rb["kept"]).all()
    c = pipe.submit(cloud)                                   # fine now
    assert pipe.collect(c)["nrows"] == ra["nrows"]
    for _ in range(5):                                       # results that stay on the device need no collect
        pipe.submit(device_cloud=torch.from_numpy(cloud).cuda(), readback=False)
