"""Training path (BASELINE.json config 5): label generation (utils/label_generation.py), loss (models/loss.py) and
the training-mode forward, against golden vectors produced by importing and running the reference unmodified on the
CPU (tests/golden/make_golden.py training).  The loss is plain torch and is checked on the CPU as well; label
generation needs the CUDA knn."""
import os

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
STAT_REL = 2e-5                       # loss terms: sums of ~1e4 fp32 terms in another order than the reference's CPU run


@pytest.fixture
def fp32_convs():
    """the composed path's cuDNN convolutions in fp32 (torch allows TF32 for them by default)"""
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32 = old


def golden(name):
    return np.load(os.path.join(HERE, "golden", "training_%s.npz" % name))


def label_inputs(b200, g, device):
    seed = int(g["seed"])
    b, n, ns, nobj, npts = [int(x) for x in g["dims"]]
    cloud = np.stack([b200.synth.tabletop_cloud(n, seed + i) for i in range(b)])
    ep = dict(b200.synth.training_labels(cloud, num_objects=nobj, num_grasp_points=npts, seed=seed, device=device))
    inds = g["fp2_inds"]
    ep["input_xyz"] = torch.from_numpy(cloud).to(device)
    ep["fp2_inds"] = torch.from_numpy(inds).to(device)
    ep["fp2_xyz"] = torch.from_numpy(np.stack([cloud[i][inds[i]] for i in range(b)])).to(device)
    ep.update({k: torch.from_numpy(v).to(device) for k, v in b200.synth.training_predictions(seed + 200, b, ns).items()})
    return ep


def check_stats(ep, g):
    for k in g.files:
        if k.startswith("stat:"):
            want, got = float(g[k]), float(ep[k[5:]])
            assert abs(got - want) <= STAT_REL * max(1.0, abs(want)), (k, got, want)


@pytest.mark.parametrize("name", ["a", "b"])
def test_loss_equals_reference_on_cpu(b200, name):
    """get_loss on the reference's own matched labels (from the golden file) and the seeded predictions, CPU tensors"""
    g = golden(name)
    seed = int(g["seed"])
    b, n, ns, nobj, npts = [int(x) for x in g["dims"]]
    cloud = np.stack([b200.synth.tabletop_cloud(n, seed + i) for i in range(b)])
    ep = {"objectness_label": b200.synth.training_labels(cloud, num_objects=nobj, num_grasp_points=npts,
                                                         seed=seed)["objectness_label"]}
    ep["fp2_inds"] = torch.from_numpy(g["fp2_inds"])
    ep.update({k: torch.from_numpy(v) for k, v in b200.synth.training_predictions(seed + 200, b, ns).items()})
    for k in ("batch_grasp_view_label", "batch_grasp_label", "batch_grasp_offset", "batch_grasp_tolerance"):
        ep[k] = torch.from_numpy(g[k])
    loss, ep = b200.loss.get_loss(ep)
    check_stats(ep, g)
    assert abs(float(loss) - float(g["stat:loss/overall_loss"])) <= STAT_REL * float(g["stat:loss/overall_loss"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["a", "b"])
@pytest.mark.parametrize("fused", [True, False])
def test_label_generation_equals_reference(b200, name, fused):
    g = golden(name)
    ep = label_inputs(b200, g, "cuda")
    ep = b200.label_generation.process_grasp_labels(ep, fused=fused)
    assert ("_label_tables" in ep) == fused
    # the nearest grasp point of every seed: the reference's CPU knn and this knn agree exactly away from ties
    assert np.abs(ep["batch_grasp_point"].cpu().numpy() - g["batch_grasp_point"]).max() <= 1e-6
    vl = ep["batch_grasp_view_label"].cpu().numpy()
    assert vl.shape == g["batch_grasp_view_label"].shape
    assert np.abs(vl - g["batch_grasp_view_label"]).max() <= 1e-5 * g["batch_grasp_view_label"].max()
    assert ((vl == 0) == (g["batch_grasp_view_label"] == 0)).all()
    rot, labels, offsets, tol, ep = b200.label_generation.match_grasp_view_and_label(ep)
    assert np.abs(rot.cpu().numpy() - g["batch_grasp_view_rot"]).max() <= 2e-6
    assert np.abs(labels.cpu().numpy() - g["batch_grasp_label"]).max() <= 1e-5 * g["batch_grasp_label"].max()
    assert (offsets.cpu().numpy() == g["batch_grasp_offset"]).all()
    assert (tol.cpu().numpy() == g["batch_grasp_tolerance"]).all()
    loss, ep = b200.loss.get_loss(ep)
    check_stats(ep, g)


@pytest.mark.gpu
def test_fused_labels_equal_composed_at_full_view_count(b200):
    """one scene with 5 objects x 300 grasp points, 1024 seeds: the two implementations agree on every output"""
    cloud = b200.synth.tabletop_cloud(20000, 3)[None]
    lab = b200.synth.training_labels(cloud, num_objects=5, num_grasp_points=300, seed=4, device="cuda")
    rng = np.random.default_rng(1)
    inds = rng.choice(20000, 1024, replace=False)
    outs = []
    for fused in (True, False):
        ep = dict(lab)
        ep["fp2_xyz"] = torch.from_numpy(cloud[:, inds]).cuda()
        ep["grasp_top_view_inds"] = torch.from_numpy(rng.integers(0, 300, (1, 1024))).cuda() if not outs else outs[0][5]
        ep = b200.label_generation.process_grasp_labels(ep, fused=fused)
        top = ep["grasp_top_view_inds"]
        rot, labels, offsets, tol, ep = b200.label_generation.match_grasp_view_and_label(ep)
        outs.append((ep["batch_grasp_point"], ep["batch_grasp_view_label"], rot, labels, offsets, top, tol))
    f, c = outs
    assert (f[0] - c[0]).abs().max() <= 1e-6                              # batched einsum vs per-object matmul: last ulp
    assert torch.equal(f[4], c[4]) and torch.equal(f[6], c[6])            # same nearest points, same rows
    assert (f[1] - c[1]).abs().max() <= 2e-6 * c[1].abs().max()           # logf (CUDA libm) vs torch.log
    assert (f[3] - c[3]).abs().max() <= 2e-6 * c[3].abs().max()
    assert (f[2] - c[2]).abs().max() <= 2e-6


@pytest.mark.gpu
def test_training_forward_equals_reference(b200, fp32_convs):
    """GraspNet(is_training=True).train(): the composed path (this package's ops with autograd + torch layers,
    BatchNorm on batch statistics), label generation and loss against the reference's own training forward"""
    g = golden("forward")
    seed = int(g["seed"])
    b, n = [int(x) for x in g["dims"]]
    net = b200.graspnet.GraspNet(is_training=True)
    b200.synth.fill_state_dict(net, seed)
    net = net.cuda().train()
    cloud = np.stack([b200.synth.tabletop_cloud(n, seed + i) for i in range(b)])
    ep = dict(b200.synth.training_labels(cloud, num_objects=3, num_grasp_points=40, seed=seed, device="cuda"))
    ep["point_clouds"] = torch.from_numpy(cloud).cuda()
    ep = net(ep)
    loss, ep = b200.loss.get_loss(ep)
    assert (ep["fp2_inds"].cpu().numpy() == g["fp2_inds"]).all()
    assert np.abs(ep["objectness_score"].detach().cpu().numpy() - g["objectness_score"]).max() <= 1e-4 * np.abs(g["objectness_score"]).max()
    same_view = (ep["grasp_top_view_inds"].cpu().numpy() == g["grasp_top_view_inds"]).mean()
    assert same_view >= 0.995
    for k in g.files:
        if k.startswith("stat:"):
            want, got = float(g[k]), float(ep[k[5:]].detach())
            assert abs(got - want) <= 2e-3 * max(1.0, abs(want)), (k, got, want)   # a few seeds may pick another top view
    loss.backward()
    grads = [p.grad for p in net.parameters()]
    assert all(gr is not None and torch.isfinite(gr).all() for gr in grads)
    assert sum(float(gr.abs().sum()) for gr in grads) > 0


@pytest.mark.gpu
def test_training_gradients_equal_legacy_ops(b200, legacy, fp32_convs):
    """every parameter's gradient of the whole training step (backbone, ApproachNet, CloudCrop, heads) with this
    library's ops against the same step with the reference's legacy CUDA extension answering the pointnet2 ops"""
    if legacy is None:
        pytest.skip("reference extensions not built (oracle/_ref)")
    from b200grasp import pointnet2_utils
    seed, b, n = 23, 2, 6000
    cloud = np.stack([b200.synth.tabletop_cloud(n, seed + i) for i in range(b)])
    labels = b200.synth.training_labels(cloud, num_objects=3, num_grasp_points=60, seed=seed, device="cuda")

    def step(ext):
        torch.manual_seed(0)
        net = b200.graspnet.GraspNet(is_training=True)
        b200.synth.fill_state_dict(net, seed)
        net = net.cuda().train()
        old = pointnet2_utils._ext
        if ext is not None:
            pointnet2_utils._ext = ext
        try:
            ep = dict(labels)
            ep["point_clouds"] = torch.from_numpy(cloud).cuda()
            ep = net(ep)
            loss, ep = b200.loss.get_loss(ep)
            loss.backward()
        finally:
            pointnet2_utils._ext = old
        return float(loss), {k: p.grad.clone() for k, p in net.named_parameters()}

    l_ours, g_ours = step(None)
    l_leg, g_leg = step(legacy["ext"])
    assert abs(l_ours - l_leg) <= 1e-5 * abs(l_leg)
    top = max(float(g.abs().max()) for g in g_leg.values())
    for k in g_leg:
        # (a conv bias in front of a BatchNorm has a zero gradient up to rounding: measure those against the step's scale)
        scale = max(float(g_leg[k].abs().max()), 1e-4 * top)
        assert float((g_ours[k] - g_leg[k]).abs().max()) <= 2e-4 * scale, k          # atomics order differs
