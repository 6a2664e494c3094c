"""The pointnet2 / knn ops as torch custom operators: torch.ops.b200grasp.<op> (north_star: "a thin C-ABI torch
custom-op layer").  Each operator is the C-ABI call of ops.py registered with torch.library --
schema inferred from the annotations, a FAKE implementation (output shapes / dtypes from input shapes, no kernel)
so that FakeTensor tracing, torch.compile and torch.export see through it, and register_autograd for the three ops
the reference differentiates (gather_operation, grouping_operation, three_interpolate: pointnet2_utils.py:81-255).
They replace the pybind functions of pointnet2/_ext_src/src/bindings.cpp:13-26 and knn/src/vision.cpp:3-5 one to
one; the autograd.Function drop-ins of pointnet2_utils.py keep the reference's class structure and call the same
C ABI.  CUDA only: there is no CPU implementation to register.

    import b200grasp
    idx = torch.ops.b200grasp.furthest_point_sampling(xyz, 1024)
"""
from typing import Tuple

import torch
from torch import Tensor

from . import ops

_lib = torch.library


@_lib.custom_op("b200grasp::furthest_point_sampling", mutates_args=(), device_types="cuda")
def furthest_point_sampling(points: Tensor, nsamples: int) -> Tensor:
    return ops.furthest_point_sampling(points, nsamples)


@furthest_point_sampling.register_fake
def _(points, nsamples):
    return points.new_empty((points.shape[0], nsamples), dtype=torch.int32)


@_lib.custom_op("b200grasp::gather_points", mutates_args=(), device_types="cuda")
def gather_points(points: Tensor, idx: Tensor) -> Tensor:
    return ops.gather_points(points, idx)


@gather_points.register_fake
def _(points, idx):
    return points.new_empty((points.shape[0], points.shape[1], idx.shape[1]))


@_lib.custom_op("b200grasp::gather_points_grad", mutates_args=(), device_types="cuda")
def gather_points_grad(grad_out: Tensor, idx: Tensor, n: int) -> Tensor:
    return ops.gather_points_grad(grad_out, idx, n)


@gather_points_grad.register_fake
def _(grad_out, idx, n):
    return grad_out.new_empty((grad_out.shape[0], grad_out.shape[1], n))


def _gather_setup(ctx, inputs, output):
    points, idx = inputs
    ctx.save_for_backward(idx)
    ctx.n = points.shape[2]


def _gather_backward(ctx, grad):
    (idx,) = ctx.saved_tensors
    return gather_points_grad(grad.contiguous(), idx, ctx.n), None


gather_points.register_autograd(_gather_backward, setup_context=_gather_setup)


@_lib.custom_op("b200grasp::ball_query", mutates_args=(), device_types="cuda")
def ball_query(new_xyz: Tensor, xyz: Tensor, radius: float, nsample: int) -> Tensor:
    return ops.ball_query(new_xyz, xyz, radius, nsample)


@ball_query.register_fake
def _(new_xyz, xyz, radius, nsample):
    return new_xyz.new_empty((new_xyz.shape[0], new_xyz.shape[1], nsample), dtype=torch.int32)


@_lib.custom_op("b200grasp::cylinder_query", mutates_args=(), device_types="cuda")
def cylinder_query(new_xyz: Tensor, xyz: Tensor, rot: Tensor, radius: float, hmin: float, hmax: float,
                   nsample: int) -> Tensor:
    return ops.cylinder_query(new_xyz, xyz, rot, radius, hmin, hmax, nsample)


@cylinder_query.register_fake
def _(new_xyz, xyz, rot, radius, hmin, hmax, nsample):
    return new_xyz.new_empty((new_xyz.shape[0], new_xyz.shape[1], nsample), dtype=torch.int32)


@_lib.custom_op("b200grasp::group_points", mutates_args=(), device_types="cuda")
def group_points(points: Tensor, idx: Tensor) -> Tensor:
    return ops.group_points(points, idx)


@group_points.register_fake
def _(points, idx):
    return points.new_empty((points.shape[0], points.shape[1], idx.shape[1], idx.shape[2]))


@_lib.custom_op("b200grasp::group_points_grad", mutates_args=(), device_types="cuda")
def group_points_grad(grad_out: Tensor, idx: Tensor, n: int) -> Tensor:
    return ops.group_points_grad(grad_out, idx, n)


@group_points_grad.register_fake
def _(grad_out, idx, n):
    return grad_out.new_empty((grad_out.shape[0], grad_out.shape[1], n))


def _group_setup(ctx, inputs, output):
    points, idx = inputs
    ctx.save_for_backward(idx)
    ctx.n = points.shape[2]


def _group_backward(ctx, grad):
    (idx,) = ctx.saved_tensors
    return group_points_grad(grad.contiguous(), idx, ctx.n), None


group_points.register_autograd(_group_backward, setup_context=_group_setup)


@_lib.custom_op("b200grasp::three_nn", mutates_args=(), device_types="cuda")
def three_nn(unknowns: Tensor, knows: Tensor) -> Tuple[Tensor, Tensor]:
    dist2, idx = ops.three_nn(unknowns, knows)
    return dist2, idx


@three_nn.register_fake
def _(unknowns, knows):
    shape = (unknowns.shape[0], unknowns.shape[1], 3)
    return unknowns.new_empty(shape), unknowns.new_empty(shape, dtype=torch.int32)


@_lib.custom_op("b200grasp::three_interpolate", mutates_args=(), device_types="cuda")
def three_interpolate(points: Tensor, idx: Tensor, weight: Tensor) -> Tensor:
    return ops.three_interpolate(points, idx, weight)


@three_interpolate.register_fake
def _(points, idx, weight):
    return points.new_empty((points.shape[0], points.shape[1], idx.shape[1]))


@_lib.custom_op("b200grasp::three_interpolate_grad", mutates_args=(), device_types="cuda")
def three_interpolate_grad(grad_out: Tensor, idx: Tensor, weight: Tensor, m: int) -> Tensor:
    return ops.three_interpolate_grad(grad_out, idx, weight, m)


@three_interpolate_grad.register_fake
def _(grad_out, idx, weight, m):
    return grad_out.new_empty((grad_out.shape[0], grad_out.shape[1], m))


def _interp_setup(ctx, inputs, output):
    points, idx, weight = inputs
    ctx.save_for_backward(idx, weight)
    ctx.m = points.shape[2]


def _interp_backward(ctx, grad):
    idx, weight = ctx.saved_tensors
    return three_interpolate_grad(grad.contiguous(), idx, weight, ctx.m), None, None


three_interpolate.register_autograd(_interp_backward, setup_context=_interp_setup)


@_lib.custom_op("b200grasp::knn", mutates_args=(), device_types="cuda")
def knn(ref: Tensor, query: Tensor, k: int) -> Tensor:
    """-> (B, k, Nq) int64, 1-based (knn_modules.knn allocates idx and calls the extension: knn_modules.py:9-17)"""
    idx = torch.empty((ref.shape[0], k, query.shape[2]), dtype=torch.int64, device=ref.device)
    ops.knn(ref, query, idx)
    return idx


@knn.register_fake
def _(ref, query, k):
    return ref.new_empty((ref.shape[0], k, query.shape[2]), dtype=torch.int64)


OPS = ("furthest_point_sampling", "gather_points", "gather_points_grad", "ball_query", "cylinder_query", "group_points",
       "group_points_grad", "three_nn", "three_interpolate", "three_interpolate_grad", "knn")
