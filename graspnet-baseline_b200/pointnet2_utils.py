"""Drop-in for the reference's pointnet2/pointnet2_utils.py (same public names,
argument order and return conventions), backed by libb200grasp.so.

Mirrors: furthest_point_sample (:51-78), gather_operation (:81-115), three_nn
(:118-147), three_interpolate (:150-204), grouping_operation (:207-255),
ball_query (:258-287), QueryAndGroup (:290-372), GroupAll (:375-421),
cylinder_query (:424-458), CylinderQueryAndGroup (:461-555), RandomDropout (:40-48).
Additions (not in the reference): cylinder_query_multi and
CylinderQueryAndGroup.forward_multi, the one-scan CloudCrop fast path.
"""
import torch
import torch.nn as nn
from torch.autograd import Function

from . import ops as _ext


class RandomDropout(nn.Module):
    """Whole-channel feature dropout with a random rate in [0, p), no rescaling.
    (The reference class calls pt_utils.feature_dropout_no_scaling, which its own
    pytorch_utils.py does not define; GraspNet never instantiates it.)"""

    def __init__(self, p=0.5, inplace=False):
        super().__init__()
        self.p = p
        self.inplace = inplace

    def forward(self, X):
        if not self.training:
            return X
        theta = torch.Tensor(1).uniform_(0, self.p)[0].item()
        keep = (torch.rand(X.shape[0], X.shape[1], 1, device=X.device) >= theta).to(X.dtype)
        return X.mul_(keep) if self.inplace else X * keep


class FurthestPointSampling(Function):
    @staticmethod
    def forward(ctx, xyz, npoint):
        """xyz (B,N,3) -> (B,npoint) int32 indices of the furthest-point set."""
        out = _ext.furthest_point_sampling(xyz, npoint)
        ctx.mark_non_differentiable(out)
        return out

    @staticmethod
    def backward(ctx, a=None):
        return None, None


furthest_point_sample = FurthestPointSampling.apply


class GatherOperation(Function):
    @staticmethod
    def forward(ctx, features, idx):
        """features (B,C,N), idx (B,npoint) int32 -> (B,C,npoint)."""
        _, C, N = features.size()
        ctx.for_backwards = (idx, C, N)
        return _ext.gather_points(features, idx)

    @staticmethod
    def backward(ctx, grad_out):
        idx, C, N = ctx.for_backwards
        return _ext.gather_points_grad(grad_out.contiguous(), idx, N), None


gather_operation = GatherOperation.apply


class ThreeNN(Function):
    @staticmethod
    def forward(ctx, unknown, known):
        """unknown (B,n,3), known (B,m,3) -> (dist (B,n,3) L2 distances, idx (B,n,3) int32)."""
        dist2, idx = _ext.three_nn(unknown, known)
        dist = torch.sqrt(dist2)
        ctx.mark_non_differentiable(dist, idx)
        return dist, idx

    @staticmethod
    def backward(ctx, a=None, b=None):
        return None, None


three_nn = ThreeNN.apply


class ThreeInterpolate(Function):
    @staticmethod
    def forward(ctx, features, idx, weight):
        """features (B,c,m), idx (B,n,3) int32, weight (B,n,3) -> (B,c,n)."""
        m = features.size(2)
        ctx.three_interpolate_for_backward = (idx, weight, m)
        return _ext.three_interpolate(features, idx, weight)

    @staticmethod
    def backward(ctx, grad_out):
        idx, weight, m = ctx.three_interpolate_for_backward
        return _ext.three_interpolate_grad(grad_out.contiguous(), idx, weight, m), None, None


three_interpolate = ThreeInterpolate.apply


class GroupingOperation(Function):
    @staticmethod
    def forward(ctx, features, idx):
        """features (B,C,N), idx (B,npoint,nsample) int32 -> (B,C,npoint,nsample)."""
        N = features.size(2)
        ctx.for_backwards = (idx, N)
        return _ext.group_points(features, idx)

    @staticmethod
    def backward(ctx, grad_out):
        idx, N = ctx.for_backwards
        return _ext.group_points_grad(grad_out.contiguous(), idx, N), None


grouping_operation = GroupingOperation.apply


class BallQuery(Function):
    @staticmethod
    def forward(ctx, radius, nsample, xyz, new_xyz):
        """xyz (B,N,3), new_xyz (B,npoint,3) -> (B,npoint,nsample) int32."""
        out = _ext.ball_query(new_xyz, xyz, radius, nsample)
        ctx.mark_non_differentiable(out)
        return out

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None, None


ball_query = BallQuery.apply


class CylinderQuery(Function):
    @staticmethod
    def forward(ctx, radius, hmin, hmax, nsample, xyz, new_xyz, rot):
        """xyz (B,N,3), new_xyz (B,npoint,3), rot (B,npoint,9) -> (B,npoint,nsample) int32."""
        out = _ext.cylinder_query(new_xyz, xyz, rot, radius, hmin, hmax, nsample)
        ctx.mark_non_differentiable(out)
        return out

    @staticmethod
    def backward(ctx, a=None):
        return None, None, None, None, None, None, None


cylinder_query = CylinderQuery.apply


def cylinder_query_multi(radius, hmin, hmax_list, nsample, xyz, new_xyz, rot):
    """All hmax lists from ONE scan: -> (len(hmax_list),B,npoint,nsample) int32; slice h is
    bit-identical to cylinder_query(radius, hmin, hmax_list[h], nsample, xyz, new_xyz, rot)."""
    with torch.no_grad():
        return _ext.cylinder_query_multi(new_xyz, xyz, rot, radius, hmin, list(hmax_list), nsample)


def _uniform_resample(idx, nsample):
    """sample_uniformly branch (pointnet2_utils.py:332-341): per region keep the unique
    indices and pad with a random re-draw of them.  Host-side loop like the reference's."""
    unique_cnt = torch.zeros((idx.shape[0], idx.shape[1]))
    for b in range(idx.shape[0]):
        for r in range(idx.shape[1]):
            u = torch.unique(idx[b, r, :])
            k = u.shape[0]
            unique_cnt[b, r] = k
            pick = torch.randint(0, k, (nsample - k,), dtype=torch.long)
            idx[b, r, :] = torch.cat((u, u[pick]))
    return unique_cnt


def _assemble(grouped_xyz, features, idx, use_xyz):
    if features is not None:
        grouped_features = grouping_operation(features, idx)
        return torch.cat([grouped_xyz, grouped_features], dim=1) if use_xyz else grouped_features
    assert use_xyz, "Cannot have not features and not use xyz as a feature!"
    return grouped_xyz


class QueryAndGroup(nn.Module):
    """Ball query of `radius` + grouping -> (B, 3+C, npoint, nsample)."""

    def __init__(self, radius, nsample, use_xyz=True, ret_grouped_xyz=False, normalize_xyz=False,
                 sample_uniformly=False, ret_unique_cnt=False):
        super().__init__()
        self.radius, self.nsample, self.use_xyz = radius, nsample, use_xyz
        self.ret_grouped_xyz = ret_grouped_xyz
        self.normalize_xyz = normalize_xyz
        self.sample_uniformly = sample_uniformly
        self.ret_unique_cnt = ret_unique_cnt
        if self.ret_unique_cnt:
            assert self.sample_uniformly

    def forward(self, xyz, new_xyz, features=None):
        idx = ball_query(self.radius, self.nsample, xyz, new_xyz)
        if self.sample_uniformly:
            unique_cnt = _uniform_resample(idx, self.nsample)
        xyz_trans = xyz.transpose(1, 2).contiguous()
        grouped_xyz = grouping_operation(xyz_trans, idx)
        grouped_xyz -= new_xyz.transpose(1, 2).unsqueeze(-1)
        if self.normalize_xyz:
            grouped_xyz /= self.radius
        ret = [_assemble(grouped_xyz, features, idx, self.use_xyz)]
        if self.ret_grouped_xyz:
            ret.append(grouped_xyz)
        if self.ret_unique_cnt:
            ret.append(unique_cnt)
        return ret[0] if len(ret) == 1 else tuple(ret)


class GroupAll(nn.Module):
    """Groups all features -> (B, 3+C, 1, N)."""

    def __init__(self, use_xyz=True, ret_grouped_xyz=False):
        super().__init__()
        self.use_xyz = use_xyz
        self.ret_grouped_xyz = ret_grouped_xyz   # the reference forgets to store this (:383-386)

    def forward(self, xyz, new_xyz, features=None):
        grouped_xyz = xyz.transpose(1, 2).unsqueeze(2)
        if features is not None:
            grouped_features = features.unsqueeze(2)
            new_features = torch.cat([grouped_xyz, grouped_features], dim=1) if self.use_xyz else grouped_features
        else:
            new_features = grouped_xyz
        return (new_features, grouped_xyz) if self.ret_grouped_xyz else new_features


class CylinderQueryAndGroup(nn.Module):
    """Cylinder query (radius, [hmin,hmax] along the rotated x axis) + grouping,
    grouped xyz expressed in the cylinder frame when rotate_xyz."""

    def __init__(self, radius, hmin, hmax, nsample, use_xyz=True, ret_grouped_xyz=False, normalize_xyz=False,
                 rotate_xyz=True, sample_uniformly=False, ret_unique_cnt=False):
        super().__init__()
        self.radius, self.nsample, self.hmin, self.hmax = radius, nsample, hmin, hmax
        self.use_xyz = use_xyz
        self.ret_grouped_xyz = ret_grouped_xyz
        self.normalize_xyz = normalize_xyz
        self.rotate_xyz = rotate_xyz
        self.sample_uniformly = sample_uniformly
        self.ret_unique_cnt = ret_unique_cnt
        if self.ret_unique_cnt:
            assert self.sample_uniformly

    def _group(self, idx, xyz, new_xyz, rot, features):
        if self.sample_uniformly:
            unique_cnt = _uniform_resample(idx, self.nsample)
        xyz_trans = xyz.transpose(1, 2).contiguous()
        grouped_xyz = grouping_operation(xyz_trans, idx)
        grouped_xyz -= new_xyz.transpose(1, 2).unsqueeze(-1)
        if self.normalize_xyz:
            grouped_xyz /= self.radius
        if self.rotate_xyz:
            g = grouped_xyz.permute(0, 2, 3, 1).contiguous()
            g = torch.matmul(g, rot)
            grouped_xyz = g.permute(0, 3, 1, 2).contiguous()
        ret = [_assemble(grouped_xyz, features, idx, self.use_xyz)]
        if self.ret_grouped_xyz:
            ret.append(grouped_xyz)
        if self.ret_unique_cnt:
            ret.append(unique_cnt)
        return ret[0] if len(ret) == 1 else tuple(ret)

    def forward(self, xyz, new_xyz, rot, features=None):
        B, npoint, _ = new_xyz.size()
        idx = cylinder_query(self.radius, self.hmin, self.hmax, self.nsample, xyz, new_xyz,
                             rot.view(B, npoint, 9))
        return self._group(idx, xyz, new_xyz, rot, features)

    def forward_with_idx(self, idx, xyz, new_xyz, rot, features=None):
        """Grouping half of forward() for an idx produced by cylinder_query_multi."""
        return self._group(idx, xyz, new_xyz, rot, features)
