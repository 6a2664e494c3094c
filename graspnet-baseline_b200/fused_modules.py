"""Set-abstraction and CloudCrop modules whose eval-mode forward is the fused tensor-core kernel
(csrc/mlp_fused.cu): neighbourhood query -> group -> shared MLP -> max-pool with no grouped or per-layer
activation tensor in HBM.

Constructor arguments, forward signatures, outputs and state_dict keys follow the reference classes
  SharedMLP               pointnet2/pytorch_utils.py:11-36   (layer{i}.conv.weight, layer{i}.bn.bn.{weight,bias,...})
  PointnetSAModuleVotes   pointnet2/pointnet2_modules.py:164-275  (mlp_module.*)
  CloudCrop               models/modules.py:80-142                 (mlps.*)
so a reference checkpoint loads unchanged.  In training mode (BatchNorm needs the batch statistics and autograd
the activations) the modules run the composed path: the same CUDA group ops followed by the torch layers.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops, pointnet2_utils

# False = always run the composed path (the reference's own op sequence: group ops + torch layers), also in eval mode
FUSED = True
# Arithmetic of the fused fp32-grade paths (modules built with passes=3): "f16x3" = csrc/mlp_f16.cu (fp16 hi | lo
# operands, exact power-of-two scaling, tcgen05 kind::f16; also the 1x1-conv stacks between the geometry ops),
# "tf32x3" = csrc/mlp_fused.cu (3xTF32) + csrc/pointwise.cu (fp32 FMAs).  passes=1 modules always run plain TF32.
ARITH = "f16x3"
# The 1x1-conv stacks of the backbone's FP modules and of ApproachNet (K = 512 / 256 / 302: up to 96 truncating
# accumulator updates per output in fp16x3) can run on the exact-fp32 kernel of csrc/pointwise.cu instead:
# measured (profiles/r02_graspnet_parity.txt, tools/pipe_probe.py): "fp32" brings the network-level error against the
# exact (float64) network from 1.2e-5 / 2.1e-5 (view scores / objectness) to 5e-6 / 1.0e-5 and the single-scene
# latency from 0.89 to 0.85 ms, at 3365 instead of 3664 scenes/s in the stream.  Parity first: "fp32" is the default;
# the two grasp heads (one launch) and all grouped MLPs stay on the tensor cores.
STACK_ARITH = "fp32"


class _BN(nn.Sequential):
    def __init__(self, channels):
        super().__init__()
        self.add_module("bn", nn.BatchNorm2d(channels))


class _ConvBNReLU(nn.Sequential):
    def __init__(self, cin, cout, bn):
        super().__init__()
        conv = nn.Conv2d(cin, cout, kernel_size=(1, 1), bias=not bn)
        nn.init.kaiming_normal_(conv.weight)
        if conv.bias is not None:
            nn.init.constant_(conv.bias, 0)
        self.add_module("conv", conv)
        if bn:
            self.add_module("bn", _BN(cout))
        self.add_module("activation", nn.ReLU(inplace=True))


class SharedMLP(nn.Sequential):
    """1x1 conv + BatchNorm2d + ReLU per layer over (B, C, npoint, nsample)."""

    def __init__(self, args, *, bn=False):
        super().__init__()
        for i in range(len(args) - 1):
            self.add_module("layer%d" % i, _ConvBNReLU(args[i], args[i + 1], bn))

    def folded(self):
        """[(w (Cout,Cin), b (Cout))] with eval-mode BatchNorm folded in: y = relu(w x + b) per layer"""
        out = []
        for layer in self:
            w = layer.conv.weight.detach()[:, :, 0, 0].float()
            b = layer.conv.bias.detach().float() if layer.conv.bias is not None else torch.zeros_like(w[:, 0])
            if hasattr(layer, "bn"):
                bn = layer.bn.bn
                scale = bn.weight.detach() / torch.sqrt(bn.running_var + bn.eps)
                w = w * scale[:, None]
                b = (b - bn.running_mean) * scale + bn.bias.detach()
            out.append((w.contiguous(), b.contiguous()))
        return out


class _FusedMLP:
    """Packed weights of a 3-layer SharedMLP for ops.group_mlp_maxpool / group_feat_mlp_maxpool; rebuilt when the
    parameters change."""

    def __init__(self, mlp):
        self.mlp, self.key, self.layers = mlp, None, None
        self.key16, self.layers16 = None, None

    def get(self):
        params = list(self.mlp.parameters()) + list(self.mlp.buffers())
        key = tuple((p.data_ptr(), p._version) for p in params)
        if key != self.key:
            (w1, b1), (w2, b2), (w3, b3) = self.mlp.folded()
            if w1.shape[1] == 3:                         # xyz only: layer 1 runs on the CUDA cores
                first = (w1, b1)
            else:                                        # [xyz | features] -> [features | xyz | zero pad to 16]
                pad = torch.zeros(w1.shape[0], 13, dtype=w1.dtype, device=w1.device)
                first = (ops.mlp_pack_weights(torch.cat([w1[:, 3:], w1[:, :3], pad], dim=1).contiguous()), b1,
                         w1.shape[0])
            self.layers = [first, (ops.mlp_pack_weights(w2), b2, w2.shape[0]),
                           (ops.mlp_pack_weights(w3), b3, w3.shape[0])]
            self.key = key
        return self.layers

    def get_f16(self):
        """(first, [F16Layer]) for ops.group_mlp_maxpool_f16 (first = (w1, b1, max row 1-norm of w1, max |b1|)) or
        (None, [three F16Layer]) for ops.group_feat_mlp_maxpool_f16"""
        params = list(self.mlp.parameters()) + list(self.mlp.buffers())
        key = tuple((p.data_ptr(), p._version) for p in params)
        if key != self.key16:
            (w1, b1), (w2, b2), (w3, b3) = self.mlp.folded()
            if w1.shape[1] == 3:
                first = (w1, b1, float(w1.abs().sum(dim=1).max()) * (1.0 + 1e-5), float(b1.abs().max()))
                self.layers16 = (first, ops.f16_stack([(w2, b2), (w3, b3)], first_k_pad=(w1.shape[0] + 31) // 32 * 32))
            else:
                c = w1.shape[1] - 3
                self.layers16 = (None, ops.f16_stack([(w1, b1), (w2, b2), (w3, b3)], first_k_pad=(c + 3 + 31) // 32 * 32,
                                                     feat_channels=c))
            self.key16 = key
        return self.layers16


def fusable(mlp, nsample):
    """shapes csrc/mlp_fused.cu covers: three layers of widths 16..256 (multiples of 16), whole lists per 128-row
    tile; input = xyz only, or xyz + C features with C a multiple of 16 and all accumulators within TMEM"""
    dims = [mlp[0].conv.in_channels] + [l.conv.out_channels for l in mlp]
    ok = (len(mlp) == 3 and nsample in (16, 32, 64, 128) and all(d % 16 == 0 and 16 <= d <= 256 for d in dims[1:]))
    if dims[0] == 3:
        return ok
    return ok and (dims[0] - 3) % 16 == 0 and 16 <= dims[0] - 3 <= 496 and sum(dims[1:]) <= 512


class PointnetSAModuleVotes(nn.Module):
    """Set abstraction: FPS -> ball query -> group -> shared MLP -> max-pool; returns (new_xyz, new_features, inds)."""

    def __init__(self, *, mlp, npoint=None, radius=None, nsample=None, bn=True, use_xyz=True, pooling="max",
                 sigma=None, normalize_xyz=False, sample_uniformly=False, ret_unique_cnt=False, passes=3):
        super().__init__()
        if npoint is None or pooling != "max" or sample_uniformly or ret_unique_cnt:
            raise NotImplementedError("b200grasp: only the configuration GraspNet uses (npoint set, max pooling)")
        self.npoint, self.radius, self.nsample = npoint, radius, nsample
        self.use_xyz, self.normalize_xyz, self.passes = use_xyz, normalize_xyz, passes
        self.grouper = pointnet2_utils.QueryAndGroup(radius, nsample, use_xyz=use_xyz, ret_grouped_xyz=True,
                                                     normalize_xyz=normalize_xyz)
        spec = list(mlp)
        if use_xyz and len(spec) > 0:
            spec[0] += 3
        self.mlp_module = SharedMLP(spec, bn=bn)
        self._fused = _FusedMLP(self.mlp_module)
        self.features_t = None

    def forward(self, xyz, features=None, inds=None, features_t=None):
        """features (B,C,N) as in the reference; `features_t` (B,N,C) may be passed as well (the channels-last copy
        the previous fused module left in its .features_t) to skip the transpose.  After a fused forward
        self.features_t holds new_features channels-last (B,npoint,C_out)."""
        if inds is None:
            inds = pointnet2_utils.furthest_point_sample(xyz, self.npoint)
        else:
            assert inds.shape[1] == self.npoint
        self.features_t = None
        if not self.fused():                             # the reference's own sequence (pointnet2_modules.py:232-272)
            xyz_flipped = xyz.transpose(1, 2).contiguous()
            new_xyz = pointnet2_utils.gather_operation(xyz_flipped, inds).transpose(1, 2).contiguous()
            grouped_features, _ = self.grouper(xyz, new_xyz, features)
            new_features = self.mlp_module(grouped_features)
            new_features = F.max_pool2d(new_features, kernel_size=[1, new_features.size(3)]).squeeze(-1)
            return new_xyz, new_features, inds
        new_xyz = ops.gather_xyz(xyz, inds, transposed=False)
        idx = pointnet2_utils.ball_query(self.radius, self.nsample, xyz, new_xyz)
        new_features, self.features_t = self.pooled(xyz, new_xyz, idx, features, features_t)
        return new_xyz, new_features, inds

    def fused(self):
        return (FUSED and not self.training and self.use_xyz and not torch.is_grad_enabled()
                and fusable(self.mlp_module, self.nsample))

    def pooled(self, xyz, new_xyz, idx, features=None, features_t=None):
        """the fused kernel on given neighbour lists idx (B,npoint,nsample): -> (new_features (B,C,npoint),
        the same channels-last (B,npoint,C))"""
        radius = self.radius if self.normalize_xyz else 0.0
        f16 = self.passes == 3 and ARITH == "f16x3"
        if features is None and features_t is None:
            if f16:
                first, layers = self._fused.get_f16()
                return ops.group_mlp_maxpool_f16(xyz, new_xyz, idx, first, layers, radius=radius, channels_last=True)
            return ops.group_mlp_maxpool(xyz, new_xyz, idx, self._fused.get(), radius=radius, passes=self.passes,
                                         channels_last=True)
        if features_t is None:
            features_t = features.transpose(1, 2).contiguous()
        if f16:
            return ops.group_feat_mlp_maxpool_f16(xyz, new_xyz, features_t, idx, self._fused.get_f16()[1],
                                                  radius=radius, channels_last=True)
        return ops.group_feat_mlp_maxpool(xyz, new_xyz, features_t, idx, self._fused.get(), radius=radius,
                                          passes=self.passes, channels_last=True)


class CloudCrop(nn.Module):
    """Cylinder crop around every seed in its approach frame, for each grasp depth; shared MLP + max-pool.
    forward(seed_xyz (B,S,3), pointcloud (B,N,3), vp_rot (B,S,3,3)) -> (B, mlps[-1], S, num_depth)."""

    def __init__(self, nsample, seed_feature_dim, cylinder_radius=0.05, hmin=-0.02,
                 hmax_list=(0.01, 0.02, 0.03, 0.04), passes=3):
        super().__init__()
        self.nsample, self.in_dim, self.cylinder_radius = nsample, seed_feature_dim, cylinder_radius
        self.hmin, self.hmax_list, self.passes = hmin, list(hmax_list), passes
        self.groupers = [pointnet2_utils.CylinderQueryAndGroup(cylinder_radius, hmin, hmax, nsample, use_xyz=True)
                         for hmax in self.hmax_list]
        self.mlps = SharedMLP([self.in_dim, 64, 128, 256], bn=True)
        self._fused = _FusedMLP(self.mlps)

    def fused(self):
        return FUSED and not self.training and not torch.is_grad_enabled() and fusable(self.mlps, self.nsample)

    def forward(self, seed_xyz, pointcloud, vp_rot, grid=None):
        B, num_seed, _, _ = vp_rot.size()
        num_depth = len(self.groupers)
        if self.fused():
            rot = vp_rot.reshape(B, num_seed, 9).contiguous()
            idx = ops.cylinder_query_multi(seed_xyz, pointcloud, rot, self.cylinder_radius, self.hmin,
                                           self.hmax_list, self.nsample, grid=grid)
            if self.passes == 3 and ARITH == "f16x3":
                first, layers = self._fused.get_f16()
                return ops.group_mlp_maxpool_f16(pointcloud, seed_xyz, idx, first, layers, rot=rot)
            return ops.group_mlp_maxpool(pointcloud, seed_xyz, idx, self._fused.get(), rot=rot, passes=self.passes)
        grouped = [g(pointcloud, seed_xyz, vp_rot) for g in self.groupers]
        grouped = torch.stack(grouped, dim=3).view(B, -1, num_seed * num_depth, self.nsample)
        vp_features = F.max_pool2d(self.mlps(grouped), kernel_size=[1, self.nsample])
        return vp_features.view(B, -1, num_seed, num_depth)
