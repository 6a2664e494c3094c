"""Device-resident form of the reference's post-processing sequence (test.py:86-96, demo_pointcloud.py:228-249):

    grasp_preds = pred_decode(end_points)                     # models/graspnet.py:77-134
    gg = GraspGroup(grasp_preds[0].detach().cpu().numpy())
    mfcdetector = ModelFreeCollisionDetector(cloud, voxel_size=...)
    gg = gg[~mfcdetector.detect(gg, approach_dist=0.05, collision_thresh=...)]
    gg = gg.nms()

The drop-in classes reproduce each call with numpy in / numpy out.  `grasp_postprocess_device` runs the same
stages back to back on the device -- voxel down-sample, collision mask, compaction, NMS, row gather -- with every
row count in device memory: no host arithmetic, no synchronisation, capturable in a CUDA graph.  The results are
the same rows the call sequence above returns (tests/test_gpu_postprocess.py).
"""
import math

import numpy as np
import torch

from . import ops as _ext


def grasp_postprocess_device(cloud, rows, nrows, voxel_size=0.01, approach_dist=0.05, collision_thresh=0.01,
                             translation_thresh=0.03, rotation_thresh=30.0 / 180.0 * math.pi, separate_calls=False):
    """cloud (N,3) f32/f64 device tensor; rows (Mmax,17) f32 with nrows (1,) i32 valid rows (e.g. one scene of
    pred_decode_device) -> (kept rows (Mmax,17) f32 in score order, nkept (1,) i32, collision mask (Mmax,) u8).
    Up to ops.POST_CHAIN_MAX_M rows the stages after the voxel grid run as one fused call (csrc/post_chain.cu);
    `separate_calls=True` forces the call-by-call sequence (identical results)."""
    scene, nvox = _ext.voxel_down_sample(cloud.double().contiguous(), voxel_size)
    if rows.shape[0] <= _ext.POST_CHAIN_MAX_M and not separate_calls:
        return _ext.grasp_post_chain(scene, nvox, rows, nrows, voxel_size, approach_dist, collision_thresh,
                                     translation_thresh, math.cos(rotation_thresh))
    mask = _ext.collision_mask_device(scene, nvox, rows, nrows, voxel_size, approach_dist, collision_thresh)
    free, nfree = _ext.compact_rows(rows, mask, nrows)
    kept, nkept = _ext.grasp_nms_device(free, nfree, translation_thresh, math.cos(rotation_thresh))
    return kept, nkept, mask


def grasp_postprocess(cloud, grasps, voxel_size=0.01, approach_dist=0.05, collision_thresh=0.01,
                      translation_thresh=0.03, rotation_thresh=30.0 / 180.0 * math.pi, device="cuda"):
    """numpy cloud (N,3) + numpy grasp rows (M,17) -> numpy kept rows (K,17): one upload, one read-back."""
    dev = torch.device(device)
    c = torch.from_numpy(np.ascontiguousarray(cloud)).to(dev, non_blocking=True)
    g = torch.from_numpy(np.ascontiguousarray(grasps, dtype=np.float32)).to(dev, non_blocking=True)
    n = torch.full((1,), g.shape[0], dtype=torch.int32, device=dev)
    kept, nkept, _ = grasp_postprocess_device(c, g, n, voxel_size, approach_dist, collision_thresh,
                                              translation_thresh, rotation_thresh)
    k = int(nkept.item())
    return kept[:k].cpu().numpy()
