"""Voxel down-sample behind ModelFreeCollisionDetector.__init__ (the reference calls open3d's
PointCloud.voxel_down_sample, utils/collision_detector.py:32-35; open3d is not available and its
arithmetic is restated: origin = min_bound - voxel/2, key = floor((p - origin)/voxel), output = fp64
mean per occupied voxel accumulated in point order, rows sorted by key -- PARITY UNPINNED, DESIGN.md
section 2).  Runs on the GPU (csrc/voxel.cu); the numpy statement of the same arithmetic lives in
oracle/post.py and is what tests compare against, bit for bit."""
import numpy as np
import torch

from . import ops as _ext


def voxel_down_sample_device(points, voxel_size, device="cuda", return_bad=False):
    """numpy / tensor (N,3) -> (scene (N,3) f64 device tensor, nvox (1,) i32 device tensor[, nbad (1,) i32]); rows
    [0, nvox) are the voxel means; nbad = input points left out (not finite / out of the 2^21-voxel range).
    No synchronisation."""
    if isinstance(points, torch.Tensor):
        p = points.to(device=device, dtype=torch.float64).reshape(-1, 3).contiguous()
    else:
        a = np.asarray(points)
        if a.dtype == np.float32:
            # half the bytes over PCIe; float -> double on the device is exact
            p = torch.from_numpy(np.ascontiguousarray(a.reshape(-1, 3))).to(device).double()
        else:
            p = torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(-1, 3))).to(device)
    return _ext.voxel_down_sample(p, voxel_size, return_bad=return_bad)


def check_voxel_input(nbad):
    """raises if the down-sample had to leave points out (open3d raises "voxel_size is too small" for the range case)"""
    bad = int(nbad.item())
    if bad:
        raise RuntimeError("b200grasp: voxel_down_sample left out %d point(s): not finite, or voxel_size is too small "
                           "for the cloud's extent (more than 2^21 voxels along an axis)" % bad)


def voxel_down_sample(points, voxel_size, device="cuda"):
    """numpy (N,3) -> numpy (Nv,3) f64, like open3d's voxel_down_sample(...).points."""
    out, nvox, nbad = voxel_down_sample_device(points, voxel_size, device, return_bad=True)
    both = torch.cat([nvox, nbad]).cpu()
    check_voxel_input(both[1])
    return out[: int(both[0])].cpu().numpy()
