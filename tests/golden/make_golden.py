#!/usr/bin/env python
"""Generates tests/golden/*.npz by RUNNING THE REFERENCE in the build container.

Needs /root/reference (read-only) and oracle/_ref (python oracle/build_ref.py);
neither exists on the GPU box, which only reads the committed .npz files.

  collision_*.npz : utils/collision_detector.py::ModelFreeCollisionDetector.detect run
                    VERBATIM (module imported from /root/reference with a stub `open3d`
                    whose voxel_down_sample is the identity -- the module only touches
                    open3d in __init__, collision_detector.py:32-35) on small seeded
                    scenes + grasps; stores inputs and every output of detect().
  knn_*.npz       : the reference's own CPU knn (knn/src/cpu/knn_cpu.cpp through
                    knn/src/knn.h:53-63, compiled unmodified into oracle/_ref) on seeded
                    inputs; stores inputs and the 1-based int64 indices.
  decode_*.npz    : models/graspnet.py::pred_decode (lines 77-134; the function body is taken from the
                    reference file with `ast` at generation time and executed with the reference's own
                    utils/loss_utils.py::batch_viewpoint_params_to_matrix) on seeded synthetic head
                    outputs; stores the inputs and the (M,17) rows it returns.
  mlp_*.npz       : pointnet2/pytorch_utils.py::SharedMLP (the reference's own class, imported; pure torch, runs on
                    the CPU in fp32) in eval mode with seeded weights and BatchNorm statistics, applied to the
                    grouped xyz tensor the QueryAndGroup / CylinderQueryAndGroup composites build
                    (pointnet2_utils.py:345-350, 523-530; the CUDA-only grouping_operation replaced by the
                    equivalent torch indexing), then F.max_pool2d over nsample and CloudCrop's view
                    (models/modules.py:138-141).  Stores the geometry inputs, the seed of the parameters
                    (tests regenerate them with synth.mlp_params) and the pooled features.
  graspnet_*.npz  : the reference's WHOLE network -- models/graspnet.py::GraspNet(is_training=False).eval() and
                    pred_decode, imported and run unmodified on the CPU in fp32 -- on a seeded cloud with seeded
                    parameters (synth.fill_state_dict, keyed by state_dict name).  The reference's pointnet2 ops are
                    CUDA-only, so the module object `pointnet2_utils._ext` (the compiled extension, imported from
                    oracle/_ref) is replaced by a shim that answers the same ten functions with the C oracle
                    (oracle/geom_oracle.c, itself pinned to the legacy CUDA kernels on the GPU box); every line of
                    Python between those ops -- groupers, SharedMLP, FP modules, ApproachNet, CloudCrop, heads,
                    decode -- is the reference's.  Stores the cloud, the parameter seed and (sub-sampled) outputs.
No single-op pointnet2 fixture can be produced here: every pointnet2 op of the reference is
CUDA-only and this container has no GPU -- those are pinned on the GPU box against
oracle/_ref directly (tests/test_gpu_vs_legacy.py).
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"


def _stub_open3d():
    o3d = types.ModuleType("open3d")

    class _PC:
        def __init__(self):
            self.points = None

        def voxel_down_sample(self, voxel_size):
            return self

    o3d.geometry = types.SimpleNamespace(PointCloud=_PC)
    o3d.utility = types.SimpleNamespace(Vector3dVector=lambda a: np.asarray(a, dtype=np.float64))
    sys.modules["open3d"] = o3d


class _GG:
    """Minimal stand-in exposing exactly the accessors detect() reads
    (collision_detector.py:70-74) over a (M,17) array, dtype preserved."""

    def __init__(self, arr):
        self.a = arr

    translations = property(lambda s: s.a[:, 13:16])
    rotation_matrices = property(lambda s: s.a[:, 4:13].reshape(-1, 3, 3))
    heights = property(lambda s: s.a[:, 2])
    depths = property(lambda s: s.a[:, 3])
    widths = property(lambda s: s.a[:, 1])


def synth_scene(rng, n):
    """table plane + a few blobs, camera frame, metres, fp64."""
    npl = n * 6 // 10
    plane = np.stack([rng.uniform(-0.3, 0.3, npl), rng.uniform(-0.3, 0.3, npl),
                      0.6 + rng.normal(0, 0.001, npl)], 1)
    blobs = []
    for _ in range(8):
        c = np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.2, 0.2), rng.uniform(0.5, 0.58)])
        blobs.append(c + rng.uniform(-0.04, 0.04, ((n - npl) // 8 + 1, 3)))
    return np.concatenate([plane] + blobs)[:n]


def rot_from_view(towards, angle):
    """orthonormal frame whose x axis is `towards`, rotated in-plane by `angle`
    (own construction; only needs to yield realistic proper rotations)."""
    ax = towards / np.linalg.norm(towards)
    ay = np.array([-ax[1], ax[0], 0.0])
    if np.linalg.norm(ay) < 1e-9:
        ay = np.array([0.0, 1.0, 0.0])
    ay /= np.linalg.norm(ay)
    az = np.cross(ax, ay)
    R1 = np.array([[1, 0, 0], [0, np.cos(angle), -np.sin(angle)], [0, np.sin(angle), np.cos(angle)]])
    return np.stack([ax, ay, az], 1) @ R1


def synth_grasps(rng, scene, m, dtype=np.float32):
    g = np.zeros((m, 17), np.float64)
    g[:, 0] = rng.uniform(0, 1, m)
    g[:, 1] = rng.uniform(0, 0.1, m)
    g[:, 2] = 0.02
    g[:, 3] = rng.integers(1, 5, m) * 0.01
    for i in range(m):
        v = rng.normal(size=3)
        v[2] = abs(v[2]) + 0.3
        g[i, 4:13] = rot_from_view(v, rng.integers(0, 12) * np.pi / 12).reshape(-1)
    g[:, 13:16] = scene[rng.choice(len(scene), m, replace=len(scene) < m)]
    g[:, 16] = -1
    return g.astype(dtype)


def make_collision():
    _stub_open3d()
    sys.path.insert(0, os.path.join(REF, "utils"))
    import collision_detector as cd
    cases = [("a", 0, 3000, 200, np.float32, 0.01, 0.05, 0.01),
             ("b", 1, 1500, 64, np.float32, 0.005, 0.03, 0.05),
             ("c", 2, 800, 33, np.float64, 0.01, 0.005, 0.01),   # approach_dist < finger_width
             ("d", 3, 1, 5, np.float32, 0.01, 0.05, 0.01)]       # single scene point
    for name, seed, n, m, dt, voxel, ad, thr in cases:
        rng = np.random.default_rng(seed)
        scene = synth_scene(rng, max(n, 16))[:n]
        grasps = synth_grasps(rng, scene, m, dt)
        det = cd.ModelFreeCollisionDetector(scene, voxel_size=voxel)
        mask, empty, ious = det.detect(_GG(grasps), approach_dist=ad, collision_thresh=thr,
                                       return_empty_grasp=True, empty_thresh=0.01, return_ious=True)
        mask_only = det.detect(_GG(grasps), approach_dist=ad, collision_thresh=thr)
        assert (mask_only == mask).all()
        np.savez_compressed(os.path.join(HERE, f"collision_{name}.npz"), scene=scene, grasps=grasps,
                            voxel_size=voxel, approach_dist=ad, collision_thresh=thr,
                            empty_thresh=0.01, mask=mask, empty=empty, ious=np.stack(ious))
        print("collision", name, scene.shape, grasps.shape, grasps.dtype, int(mask.sum()), int(empty.sum()))


def make_knn():
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
    from knn_pytorch import knn_pytorch
    cases = [("a", 0, 1, 3, 300, 300, 1), ("b", 1, 2, 3, 700, 128, 1), ("c", 2, 1, 3, 50, 7, 4),
             ("d", 3, 1, 5, 64, 33, 3), ("e", 4, 1, 3, 200, 64, 1)]
    for name, seed, b, d, nr, nq, k in cases:
        g = torch.Generator().manual_seed(seed)
        ref = torch.randn(b, d, nr, generator=g)
        query = torch.randn(b, d, nq, generator=g)
        if name == "e":  # grid-snapped: exact distance ties
            ref = torch.round(ref * 2) / 2
            query = torch.round(query * 2) / 2
        inds = torch.empty(b, k, nq).long()
        knn_pytorch.knn(ref.contiguous(), query.contiguous(), inds)
        np.savez_compressed(os.path.join(HERE, f"knn_{name}.npz"), ref=ref.numpy(), query=query.numpy(),
                            k=k, inds=inds.numpy())
        print("knn", name, tuple(ref.shape), tuple(query.shape), k)


def _reference_pred_decode():
    """pred_decode exactly as written in the reference, without importing models/graspnet.py (that module
    imports the CUDA-only pointnet2 extension at the top)."""
    import ast
    import torch
    sys.path.insert(0, os.path.join(REF, "utils"))
    import loss_utils
    src = open(os.path.join(REF, "models", "graspnet.py")).read()
    fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "pred_decode")
    glob = {"torch": torch, "np": np, "GRASP_MAX_WIDTH": loss_utils.GRASP_MAX_WIDTH,
            "GRASP_MAX_TOLERANCE": loss_utils.GRASP_MAX_TOLERANCE,
            "batch_viewpoint_params_to_matrix": loss_utils.batch_viewpoint_params_to_matrix}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "graspnet.py:pred_decode", "exec"), glob)
    return glob["pred_decode"]


def synth_heads(seed, b, ns, a=12, d=4, ties=False):
    """head outputs with the shapes models/modules.py produces (OperationNet / ToleranceNet / ApproachNet)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    ep = {"point_clouds": torch.zeros(b, 1, 3),
          "objectness_score": torch.randn(b, 2, ns, generator=g),
          "grasp_score_pred": torch.rand(b, a, ns, d, generator=g) * 1.2 - 0.1,
          "grasp_angle_cls_pred": torch.randn(b, a, ns, d, generator=g),
          "grasp_width_pred": torch.rand(b, a, ns, d, generator=g) * 0.12 - 0.01,
          "grasp_tolerance_pred": torch.rand(b, a, ns, d, generator=g) * 0.06,
          "fp2_xyz": torch.rand(b, ns, 3, generator=g) * 0.6 - 0.3,
          "grasp_top_view_xyz": torch.nn.functional.normalize(torch.randn(b, ns, 3, generator=g), dim=-1)}
    if ties:   # equal class scores / equal depth scores / objectness ties / views along +-z (axis_y degenerate)
        ep["grasp_angle_cls_pred"] = torch.round(ep["grasp_angle_cls_pred"])
        ep["grasp_score_pred"] = torch.round(ep["grasp_score_pred"] * 4) / 4
        ep["objectness_score"] = torch.round(ep["objectness_score"])
        ep["grasp_top_view_xyz"][:, ::5] = torch.tensor([0.0, 0.0, 1.0])
        ep["grasp_top_view_xyz"][:, 1::7] = torch.tensor([0.0, 0.0, -1.0])
    return ep


def make_decode():
    pred_decode = _reference_pred_decode()
    for name, seed, b, ns, ties in (("a", 0, 1, 1024, False), ("b", 1, 2, 257, True), ("c", 2, 1, 5, True)):
        ep = synth_heads(seed, b, ns, ties=ties)
        out = pred_decode(ep)
        save = {k: v.numpy() for k, v in ep.items() if k != "point_clouds"}
        for i, o in enumerate(out):
            save["rows_%d" % i] = o.numpy()
        np.savez_compressed(os.path.join(HERE, f"decode_{name}.npz"), **save)
        print("decode", name, b, ns, [tuple(o.shape) for o in out])


def make_mlp():
    import torch
    import torch.nn.functional as F
    sys.path.insert(0, os.path.join(REF, "pointnet2"))
    import pytorch_utils as pt_utils                       # the reference's SharedMLP
    sys.path.insert(0, ROOT)
    import oracle
    oracle.cpu.build()
    sys.path.insert(0, os.path.join(ROOT, "graspnet-baseline_b200"))
    import synth
    cases = (("a", 5, [3, 64, 64, 128], 1, 600, 10, 64, 0.04, False, 1),      # SA1-like: /radius, no rotation
             ("b", 6, [3, 64, 128, 256], 2, 800, 6, 64, 0.05, True, 4),       # CloudCrop-like: 4 depth lists, rotated
             ("c", 7, [3, 32, 48, 80], 1, 300, 9, 16, 0.3, True, 2),          # odd widths, 16-row lists
             ("d", 8, [3 + 32, 48, 64, 96], 2, 400, 11, 32, 0.1, False, 1),   # SA2-like: xyz + 32 point features
             ("e", 9, [3 + 128, 128, 128, 256], 1, 256, 7, 16, 0.2, False, 1))  # SA3-like widths, 16-row lists
    for name, seed, dims, b, n, m, ns, radius, crop, g in cases:
        xyz = np.stack([synth.tabletop_cloud(n, seed + i) for i in range(b)])
        first = oracle.furthest_point_sampling(xyz, m)
        new_xyz = np.stack([xyz[i][first[i]] for i in range(b)])
        if crop:
            rot = np.stack([synth.seed_rotations(m, seed + i) for i in range(b)]).reshape(b, m, 9)
            hmax = [0.01 * (k + 1) for k in range(g)]
            idx = np.stack([oracle.cylinder_query(new_xyz, xyz, rot, radius, -0.02, h, ns) for h in hmax])
            norm_radius = 0.0
        else:
            rot = None
            idx = oracle.ball_query(new_xyz, xyz, radius, ns)[None]
            norm_radius = radius
        nfeat = dims[0] - 3
        feat = np.random.default_rng(seed).standard_normal((b, nfeat, n)).astype(np.float32) if nfeat else None
        params = synth.mlp_params(seed, dims)
        mlp = pt_utils.SharedMLP(list(dims), bn=True)
        with torch.no_grad():
            for layer, (w, gamma, beta, mean, var) in zip(mlp, params):
                layer.conv.weight.copy_(torch.from_numpy(w)[:, :, None, None])
                layer.bn.bn.weight.copy_(torch.from_numpy(gamma))
                layer.bn.bn.bias.copy_(torch.from_numpy(beta))
                layer.bn.bn.running_mean.copy_(torch.from_numpy(mean))
                layer.bn.bn.running_var.copy_(torch.from_numpy(var))
        mlp.eval()
        feats = []
        tx, tc = torch.from_numpy(xyz), torch.from_numpy(new_xyz)
        for k in range(idx.shape[0]):
            ti = torch.from_numpy(idx[k]).long()
            grouped = torch.stack([tx[i].t()[:, ti[i]] for i in range(b)])          # (B,3,m,ns) = grouping_operation
            grouped = grouped - tc.transpose(1, 2).unsqueeze(-1)
            if norm_radius:
                grouped = grouped / norm_radius
            if rot is not None:
                g_ = grouped.permute(0, 2, 3, 1).contiguous()
                g_ = torch.matmul(g_, torch.from_numpy(rot).view(b, m, 3, 3))
                grouped = g_.permute(0, 3, 1, 2).contiguous()
            if feat is not None:                                                    # pointnet2_utils.py:352-358
                tf = torch.from_numpy(feat)
                gfeat = torch.stack([tf[i][:, ti[i]] for i in range(b)])            # (B,C,m,ns) = grouping_operation
                grouped = torch.cat([grouped, gfeat], dim=1)
            feats.append(grouped)
        x = torch.stack(feats, dim=3).view(b, dims[0], m * idx.shape[0], ns)          # modules.py:131-132
        with torch.no_grad():
            y = F.max_pool2d(mlp(x), kernel_size=[1, ns]).view(b, -1, m, idx.shape[0])
        np.savez_compressed(os.path.join(HERE, f"mlp_{name}.npz"), xyz=xyz, new_xyz=new_xyz, idx=idx,
                            rot=np.zeros(0, np.float32) if rot is None else rot, radius=np.float32(norm_radius),
                            seed=seed, dims=np.array(dims), out=y.numpy(),
                            feat=np.zeros(0, np.float32) if feat is None else feat)
        print("mlp", name, dims, tuple(y.shape), float(y.abs().max()))


class _OracleExt:
    """the ten functions of pointnet2._ext (bindings.cpp:13-26) on CPU tensors, answered by the C oracle"""

    def __init__(self, oracle, torch):
        self.o, self.t = oracle, torch

    def _np(self, *ts):
        return [np.ascontiguousarray(t.detach().cpu().numpy()) for t in ts]

    def furthest_point_sampling(self, xyz, npoint):
        return self.t.from_numpy(self.o.furthest_point_sampling(*self._np(xyz), npoint))

    def gather_points(self, features, idx):
        return self.t.from_numpy(self.o.gather_points(*self._np(features, idx)))

    def ball_query(self, new_xyz, xyz, radius, nsample):
        return self.t.from_numpy(self.o.ball_query(*self._np(new_xyz, xyz), radius, nsample))

    def cylinder_query(self, new_xyz, xyz, rot, radius, hmin, hmax, nsample):
        return self.t.from_numpy(self.o.cylinder_query(*self._np(new_xyz, xyz, rot), radius, hmin, hmax, nsample))

    def group_points(self, features, idx):
        return self.t.from_numpy(self.o.group_points(*self._np(features, idx)))

    def three_nn(self, unknown, known):
        d2, idx = self.o.three_nn(*self._np(unknown, known))
        return self.t.from_numpy(d2), self.t.from_numpy(idx)

    def three_interpolate(self, features, idx, weight):
        return self.t.from_numpy(self.o.three_interpolate(*self._np(features, idx, weight)))


class _OracleExt64(_OracleExt):
    """the same for the float64 run of the network: every INDEX op sees the float32 coordinates (so the indices are
    the fp32 run's), the three feature ops -- gather / group (copies) and three_interpolate (a 3-tap weighted sum) --
    are evaluated in float64 with numpy"""

    def _f32(self, t):
        return np.ascontiguousarray(t.detach().cpu().numpy().astype(np.float32))

    def furthest_point_sampling(self, xyz, npoint):
        return self.t.from_numpy(self.o.furthest_point_sampling(self._f32(xyz), npoint))

    def ball_query(self, new_xyz, xyz, radius, nsample):
        return self.t.from_numpy(self.o.ball_query(self._f32(new_xyz), self._f32(xyz), radius, nsample))

    def cylinder_query(self, new_xyz, xyz, rot, radius, hmin, hmax, nsample):
        return self.t.from_numpy(self.o.cylinder_query(self._f32(new_xyz), self._f32(xyz), self._f32(rot), radius, hmin,
                                                       hmax, nsample))

    def three_nn(self, unknown, known):
        d2, idx = self.o.three_nn(self._f32(unknown), self._f32(known))
        return self.t.from_numpy(d2.astype(np.float64)), self.t.from_numpy(idx)

    def gather_points(self, features, idx):
        f, i = features.numpy(), idx.numpy()
        return self.t.from_numpy(np.stack([f[b][:, i[b]] for b in range(len(f))]))

    def group_points(self, features, idx):
        f, i = features.numpy(), idx.numpy()
        return self.t.from_numpy(np.stack([f[b][:, i[b]] for b in range(len(f))]))

    def three_interpolate(self, features, idx, weight):
        f, i, w = features.numpy(), idx.numpy(), weight.numpy()
        return self.t.from_numpy(np.stack([(f[b][:, i[b]] * w[b][None]).sum(axis=-1) for b in range(len(f))]))


def make_graspnet():
    import torch
    sys.path.insert(0, ROOT)
    import oracle
    oracle.cpu.build()
    sys.path.insert(0, os.path.join(ROOT, "graspnet-baseline_b200"))
    import synth
    for d in ("models", "pointnet2", "utils", "knn"):
        sys.path.insert(0, os.path.join(REF, d))
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))       # pointnet2._ext, knn_pytorch (import-time only)
    import pointnet2_utils as ref_pu
    ref_pu._ext = _OracleExt(oracle, torch)
    import graspnet as ref_graspnet                                  # models/graspnet.py
    torch.set_num_threads(os.cpu_count() or 1)
    for name, seed, b, n in (("a", 63, 1, 5000), ("b", 84, 2, 3000)):
        net = ref_graspnet.GraspNet(input_feature_dim=0, num_view=300, num_angle=12, num_depth=4,
                                    cylinder_radius=0.05, hmin=-0.02, hmax_list=[0.01, 0.02, 0.03, 0.04],
                                    is_training=False)
        synth.fill_state_dict(net, seed)
        net.eval()
        cloud = np.stack([synth.tabletop_cloud(n, seed + i) for i in range(b)])
        with torch.no_grad():
            ep = net({"point_clouds": torch.from_numpy(cloud)})
            rows = ref_graspnet.pred_decode(ep)
        sd = net.state_dict()
        save = {"cloud": cloud, "seed": seed, "state_keys": np.array(sorted(sd)),
                "state_shapes": np.array([",".join(map(str, sd[k].shape)) for k in sorted(sd)]),
                "sa1_inds": ep["sa1_inds"].numpy(), "fp2_inds": ep["fp2_inds"].numpy(),
                "sa1_features": ep["sa1_features"].numpy()[:, :, ::16],
                "sa4_features": ep["sa4_features"].numpy()[:, :, ::4],
                "fp2_features": ep["fp2_features"].numpy()[:, :, ::8],
                "objectness_score": ep["objectness_score"].numpy(),
                "grasp_top_view_inds": ep["grasp_top_view_inds"].numpy(),
                "grasp_top_view_score": ep["grasp_top_view_score"].numpy(),
                "view_score": ep["view_score"].numpy()[:, ::8],
                "grasp_top_view_xyz": ep["grasp_top_view_xyz"].numpy(),
                "grasp_top_view_rot": ep["grasp_top_view_rot"].numpy()}
        for k in ("grasp_score_pred", "grasp_angle_cls_pred", "grasp_width_pred", "grasp_tolerance_pred"):
            save[k] = ep[k].numpy()[:, :, ::4]
        for i, r in enumerate(rows):
            save["rows_%d" % i] = r.numpy()
        # the same network and cloud evaluated in float64 (same indices: the index ops see the float32 coordinates):
        # the EXACT answer up to 1e-12, against which the reference's own float32 result above and any other float32
        # implementation can be measured (tests/test_gpu_graspnet.py::test_error_against_exact_network)
        ref_pu._ext = _OracleExt64(oracle, torch)
        net64 = net.double()
        with torch.no_grad():                              # models/graspnet.py:70-75, the two stages one by one: the view
            ep64 = net64.view_estimator({"point_clouds": torch.from_numpy(cloud).double()})   # frames come out in float32
            ep64["grasp_top_view_rot"] = ep64["grasp_top_view_rot"].double()
            ep64 = net64.grasp_generator(ep64)
        ref_pu._ext = _OracleExt(oracle, torch)
        assert (ep64["sa1_inds"].numpy() == ep["sa1_inds"].numpy()).all()
        assert (ep64["fp2_inds"].numpy() == ep["fp2_inds"].numpy()).all()
        for k, sl in (("sa4_features", (slice(None), slice(None), slice(None, None, 4))),
                      ("fp2_features", (slice(None), slice(None), slice(None, None, 8))),
                      ("objectness_score", (slice(None),)), ("view_score", (slice(None), slice(None, None, 8))),
                      ("grasp_top_view_score", (slice(None),))):
            save["exact:" + k] = ep64[k].numpy()[sl]
        save["exact:grasp_top_view_inds"] = ep64["grasp_top_view_inds"].numpy()
        np.savez_compressed(os.path.join(HERE, f"graspnet_{name}.npz"), **save)
        print("graspnet", name, cloud.shape, [tuple(r.shape) for r in rows],
              "fp2 |max|", float(ep["fp2_features"].abs().max()))


def make_training():
    """utils/label_generation.py (process_grasp_labels, match_grasp_view_and_label), models/loss.py (get_loss) and the
    training-mode forward of models/graspnet.py, imported and run unmodified on the CPU (knn = the reference's
    knn_cpu.cpp built into oracle/_ref, pointnet2._ext answered by the C oracle)."""
    import torch
    sys.path.insert(0, ROOT)
    import oracle
    oracle.cpu.build()
    sys.path.insert(0, os.path.join(ROOT, "graspnet-baseline_b200"))
    import synth
    for d in ("models", "pointnet2", "utils", "knn"):
        sys.path.insert(0, os.path.join(REF, d))
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
    import pointnet2_utils as ref_pu
    ref_pu._ext = _OracleExt(oracle, torch)
    import graspnet as ref_graspnet
    import label_generation as ref_lg
    import loss as ref_loss
    torch.set_num_threads(os.cpu_count() or 1)
    # ---- (1) labels + loss on seeded stand-in predictions
    for name, seed, b, n, ns, nobj, npts in (("a", 5, 2, 3000, 64, 3, 40), ("b", 9, 1, 2000, 96, 2, 25)):
        cloud = np.stack([synth.tabletop_cloud(n, seed + i) for i in range(b)])
        lab = synth.training_labels(cloud, num_objects=nobj, num_grasp_points=npts, seed=seed)
        rng = np.random.default_rng(seed + 100)
        inds = np.stack([rng.choice(n, ns, replace=False) for _ in range(b)]).astype(np.int32)
        ep = dict(lab)
        ep["input_xyz"] = torch.from_numpy(cloud)
        ep["fp2_inds"] = torch.from_numpy(inds)
        ep["fp2_xyz"] = torch.from_numpy(np.stack([cloud[i][inds[i]] for i in range(b)]))
        ep.update({k: torch.from_numpy(v) for k, v in synth.training_predictions(seed + 200, b, ns).items()})
        ep = ref_lg.process_grasp_labels(ep)
        save = {"seed": seed, "dims": np.array([b, n, ns, nobj, npts]), "fp2_inds": inds,
                "batch_grasp_point": ep["batch_grasp_point"].numpy(),
                "batch_grasp_view_label": ep["batch_grasp_view_label"].numpy(),
                "u_max_check": np.float32(0)}
        rot, labels, offsets, tol, ep = ref_lg.match_grasp_view_and_label(ep)
        save.update({"batch_grasp_view_rot": rot.numpy(), "batch_grasp_label": labels.numpy(),
                     "batch_grasp_offset": offsets.numpy(), "batch_grasp_tolerance": tol.numpy()})
        loss, ep = ref_loss.get_loss(ep)
        for k in sorted(ep):
            if k.startswith("loss/") or k.startswith("stage"):
                save["stat:" + k] = np.float64(ep[k].item())
        np.savez_compressed(os.path.join(HERE, f"training_{name}.npz"), **save)
        print("training", name, "loss", float(loss), "view label max", float(ep["batch_grasp_view_label"].max()))
    # ---- (2) the whole training-mode forward (BatchNorm on batch statistics) + loss
    seed, b, n = 17, 2, 5000
    net = ref_graspnet.GraspNet(input_feature_dim=0, num_view=300, num_angle=12, num_depth=4, cylinder_radius=0.05,
                                hmin=-0.02, hmax_list=[0.01, 0.02, 0.03, 0.04], is_training=True)
    synth.fill_state_dict(net, seed)
    net.train()
    cloud = np.stack([synth.tabletop_cloud(n, seed + i) for i in range(b)])
    ep = dict(synth.training_labels(cloud, num_objects=3, num_grasp_points=40, seed=seed))
    ep["point_clouds"] = torch.from_numpy(cloud)
    with torch.no_grad():
        ep = net(ep)
        loss, ep = ref_loss.get_loss(ep)
    save = {"seed": seed, "dims": np.array([b, n]), "fp2_inds": ep["fp2_inds"].numpy(),
            "grasp_top_view_inds": ep["grasp_top_view_inds"].numpy(),
            "objectness_score": ep["objectness_score"].numpy(),
            "grasp_score_pred": ep["grasp_score_pred"].numpy()[:, :, ::8],
            "batch_grasp_view_label": ep["batch_grasp_view_label"].numpy()[:, ::16]}
    for k in sorted(ep):
        if k.startswith("loss/"):
            save["stat:" + k] = np.float64(ep[k].item())
    np.savez_compressed(os.path.join(HERE, "training_forward.npz"), **save)
    print("training forward loss", float(loss))


def make_scene():
    """BASELINE.json configs[0]: the reference's own fixture scene doc/example_data/{depth.png, meta.mat,
    workspace_mask.png} -> point cloud with the reference's utils/data_utils.py:18-43 imported unchanged (recipe of
    image_demo.py:95-134), voxel 0.01, 1024 synthetic grasps, the reference's detect() run verbatim.  The file keeps the
    depth image (cropped to the mask's bounding box, uint16), the mask and the intrinsics -- 0.4 MB instead of 12 MB of
    points -- plus a subsample of the cloud the reference code produced, so the test can rebuild and pin it."""
    import scipy.io as scio
    from PIL import Image
    sys.path.insert(0, ROOT)
    import oracle
    oracle.cpu.build()
    sys.path.insert(0, os.path.join(ROOT, "graspnet-baseline_b200"))
    import synth
    _stub_open3d()
    sys.path.insert(0, os.path.join(REF, "utils"))
    from data_utils import CameraInfo, create_point_cloud_from_depth_image
    import collision_detector as ref_cd
    d = os.path.join(REF, "doc", "example_data")
    depth = np.array(Image.open(os.path.join(d, "depth.png")))
    meta = scio.loadmat(os.path.join(d, "meta.mat"))
    K, factor = meta["intrinsic_matrix"], float(meta["factor_depth"].item())
    wmask = np.array(Image.open(os.path.join(d, "workspace_mask.png")).convert("L"))
    cam = CameraInfo(depth.shape[1], depth.shape[0], K[0, 0], K[1, 1], K[0, 2], K[1, 2], factor)
    cloud = create_point_cloud_from_depth_image(depth, cam, organized=True)
    mask = (wmask & (depth > 0)).astype(bool)
    pts = cloud[mask]                                                   # (513688, 3) float64
    ys, xs = np.where(mask)
    y0, y1, x0, x1 = ys.min(), ys.max() + 1, xs.min(), xs.max() + 1
    # 20000-point sample as demo_pointcloud.py:178-186 draws it (without replacement when there are enough points)
    rng = np.random.RandomState(0)
    idxs = rng.choice(len(pts), 20000, replace=False)
    sample = pts[idxs].astype(np.float32)
    # the detector on the full cloud at voxel 0.01 (open3d restated: oracle.voxel_down_sample), 1024 synthetic grasps
    scene_v = oracle.voxel_down_sample(pts.astype(np.float64), 0.01)
    grasps = synth.synthetic_grasps(sample, 1024, 7)
    det = ref_cd.ModelFreeCollisionDetector.__new__(ref_cd.ModelFreeCollisionDetector)
    det.finger_width, det.finger_length, det.voxel_size, det.scene_points = 0.01, 0.06, 0.01, scene_v
    gg = _GG(grasps)
    m1, empty, ious = det.detect(gg, approach_dist=0.05, collision_thresh=0.01, return_empty_grasp=True, return_ious=True)
    det.scene_points = sample.astype(np.float64)                        # and against the 20000-point sample, unvoxelised
    m2 = det.detect(gg, approach_dist=0.05, collision_thresh=0.01)
    np.savez_compressed(os.path.join(HERE, "scene_example.npz"),
                        depth_crop=depth[y0:y1, x0:x1].astype(np.uint16), mask_crop=np.packbits(mask[y0:y1, x0:x1]),
                        crop=np.array([y0, y1, x0, x1]), shape=np.array(depth.shape), intrinsic=K.astype(np.float64),
                        factor_depth=np.float64(factor), num_points=len(pts), cloud_every_997=pts[::997],
                        cloud_sum=pts.sum(axis=0), sample_idx=idxs.astype(np.int32), num_voxels=len(scene_v),
                        grasp_seed=7, mask_voxel=m1, empty_voxel=empty, ious_voxel=np.stack(ious), mask_sample=m2)
    print("scene", pts.shape, "voxels", len(scene_v), "colliding", int(m1.sum()), int(m2.sum()),
          os.path.getsize(os.path.join(HERE, "scene_example.npz")))


if __name__ == "__main__":
    which = sys.argv[1:] or ["collision", "knn", "decode", "mlp", "graspnet", "training", "scene"]
    if "scene" in which:
        make_scene()
        sys.exit(0) if which == ["scene"] else None
    if "training" in which:
        make_training()
        sys.exit(0) if which == ["training"] else None
    if "graspnet" in which:
        make_graspnet()
        sys.exit(0) if which == ["graspnet"] else None
    if "mlp" in which:
        make_mlp()
    if "collision" in which:
        make_collision()
    if "knn" in which:
        make_knn()
    if "decode" in which:
        make_decode()
