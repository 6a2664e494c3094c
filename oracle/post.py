"""Oracle for the post-processing rows: collision detection, voxel down-sample, grasp NMS.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Grasp rows follow the 17-column layout fixed by models/graspnet.py:131-133:
[score, width, height, depth, R(9 row-major), t(3), object_id].
"""
import ctypes
import math

import numpy as np

from .cpu import lib, _p, _ci

FINGER_WIDTH = 0.01   # utils/collision_detector.py:29
FINGER_LENGTH = 0.06  # utils/collision_detector.py:30


def collision_thresholds(grasps, approach_dist):
    """Per-grasp half-space bounds with the reference's own dtype rules
    (utils/collision_detector.py:69-93): grasp fields keep the array dtype (fp32
    when the caller passes `.cpu().numpy()`, test.py:88-89), Python floats are
    weak scalars, so e.g. `depths - finger_length` is rounded to fp32 BEFORE it
    is compared with the fp64 targets.  Returns (M,10) fp64 holding those values.
    """
    approach_dist = max(approach_dist, FINGER_WIDTH)
    h = grasps[:, 2]
    d = grasps[:, 3]
    w = grasps[:, 1]
    cols = [-h / 2, h / 2, d - FINGER_LENGTH, d, -(w / 2 + FINGER_WIDTH), -w / 2,
            (w / 2 + FINGER_WIDTH), w / 2, d - FINGER_LENGTH - FINGER_WIDTH,
            d - FINGER_LENGTH - FINGER_WIDTH - approach_dist]
    return np.stack([c.astype(np.float64) for c in cols], axis=1)


def collision_counts(scene_points, grasps, approach_dist, mode=1):
    """Six integer counts per grasp (global,left,right,bottom,shifting,inner):
    collision_detector.py:75-100,119.  mode=1 is numpy's matmul association on
    FMA hardware, fma(z,r2, fma(y,r1, x*r0)) (pinned by tests/golden/collision_*.npz)."""
    scene = np.ascontiguousarray(scene_points, dtype=np.float64)
    g = np.asarray(grasps)
    T = np.ascontiguousarray(g[:, 13:16], dtype=np.float64)
    R = np.ascontiguousarray(g[:, 4:13], dtype=np.float64)
    thr = np.ascontiguousarray(collision_thresholds(g, approach_dist))
    m = g.shape[0]
    counts = np.zeros((m, 6), np.int32)
    lib().oracle_collision_counts(_p(scene), _ci(scene.shape[0]), _p(T), _p(R), _p(thr), _ci(m),
                                  _ci(mode), _p(counts))
    return counts


def collision_finalize(counts, grasps, voxel_size, approach_dist, collision_thresh,
                       return_empty_grasp=False, empty_thresh=0.01, return_ious=False):
    """collision_detector.py:103-129 from integer counts, same expressions and dtypes."""
    approach_dist = max(approach_dist, FINGER_WIDTH)
    g = np.asarray(grasps)
    heights = g[:, 2][:, np.newaxis]
    widths = g[:, 1][:, np.newaxis]
    c = counts.astype(np.int64)
    left_right_volume = (heights * FINGER_LENGTH * FINGER_WIDTH / (voxel_size ** 3)).reshape(-1)
    bottom_volume = (heights * (widths + 2 * FINGER_WIDTH) * FINGER_WIDTH / (voxel_size ** 3)).reshape(-1)
    shifting_volume = (heights * (widths + 2 * FINGER_WIDTH) * approach_dist / (voxel_size ** 3)).reshape(-1)
    volume = left_right_volume * 2 + bottom_volume + shifting_volume
    global_iou = c[:, 0] / (volume + 1e-6)
    collision_mask = (global_iou > collision_thresh)
    if not (return_empty_grasp or return_ious):
        return collision_mask
    ret = [collision_mask]
    if return_empty_grasp:
        inner_volume = (heights * FINGER_LENGTH * widths / (voxel_size ** 3)).reshape(-1)
        ret.append(c[:, 5] / inner_volume < empty_thresh)
    if return_ious:
        ret.append([global_iou, c[:, 1] / (left_right_volume + 1e-6), c[:, 2] / (left_right_volume + 1e-6),
                    c[:, 3] / (bottom_volume + 1e-6), c[:, 4] / (shifting_volume + 1e-6)])
    return ret


def collision_detect(scene_points, grasps, voxel_size, approach_dist=0.03, collision_thresh=0.05,
                     return_empty_grasp=False, empty_thresh=0.01, return_ious=False):
    """ModelFreeCollisionDetector.detect (collision_detector.py:37-129) on an
    already down-sampled fp64 scene."""
    counts = collision_counts(scene_points, grasps, approach_dist)
    return collision_finalize(counts, grasps, voxel_size, approach_dist, collision_thresh,
                              return_empty_grasp, empty_thresh, return_ious)


def voxel_down_sample(points, voxel_size, origin_at_min=False):
    """Restatement of Open3D PointCloud::VoxelDownSample, called at
    collision_detector.py:32-35 (third-party, un-vendored: open3d>=0.8,
    requirements.txt:5 -- PARITY UNPINNED).  origin = min_bound - voxel/2; voxel
    key = floor((p-origin)/voxel); output = fp64 mean of each occupied voxel,
    accumulated in point order; rows returned sorted by key (Open3D's order is
    its hash-map order, which detect() never depends on: it only counts)."""
    p = np.asarray(points, dtype=np.float64)
    if p.shape[0] == 0:
        return p.reshape(0, 3)
    # points that are not finite, or beyond 2^21 voxels per axis from the minimum corner, are left out (csrc/voxel.cu;
    # open3d raises "voxel_size is too small" for the range case)
    p = p[np.isfinite(p).all(axis=1)]
    if p.shape[0] == 0:
        return p.reshape(0, 3)
    origin = p.min(axis=0) - (0.0 if origin_at_min else voxel_size * 0.5)     # B200_VOXEL_ORIGIN_AT_MIN
    q = np.floor((p - origin) / voxel_size)
    p = p[(q < 2097152.0).all(axis=1)]
    key = np.floor((p - origin) / voxel_size).astype(np.int64)
    order = np.lexsort((key[:, 2], key[:, 1], key[:, 0]))  # stable
    ks = key[order]
    first = np.ones(len(ks), bool)
    first[1:] = np.any(ks[1:] != ks[:-1], axis=1)
    starts = np.flatnonzero(first)
    ends = np.append(starts[1:], len(ks))
    out = np.empty((len(starts), 3), np.float64)
    ps = p[order]
    for v, (s, e) in enumerate(zip(starts, ends)):  # sequential in-order accumulation
        acc = np.zeros(3)
        for r in range(s, e):
            acc = acc + ps[r]
        out[v] = acc / float(e - s)
    return out


def grasp_nms(grasps, translation_thresh=0.03, rotation_thresh=30.0 / 180.0 * math.pi, flags=0):
    """GraspGroup.nms -> grasp_nms.nms_grasp (third-party C extension, un-vendored,
    no source under /root/reference: PARITY UNPINNED; call sites
    demo_pointcloud.py:249 etc.).  Returns the kept ROW NUMBERS in score order.
    flags: the open choices of the restatement (B200_NMS_* of include/b200grasp.h); 0 = the default."""
    g = np.ascontiguousarray(grasps, dtype=np.float32)
    m = g.shape[0]
    keep = np.zeros(max(m, 1), np.int32)
    if flags:
        import ctypes
        nk = lib().oracle_grasp_nms_opts(_p(g), m, ctypes.c_double(float(translation_thresh)),
                                         ctypes.c_double(math.cos(rotation_thresh)), ctypes.c_uint(flags), _p(keep))
    else:
        nk = lib().oracle_grasp_nms(_p(g), m, float(translation_thresh), math.cos(rotation_thresh), _p(keep))
    return keep[:nk].copy()


GRASP_MAX_WIDTH = 0.1        # utils/loss_utils.py:8
GRASP_MAX_TOLERANCE = 0.05   # utils/loss_utils.py:9


def viewpoint_params_to_matrix(towards, angle):
    """utils/loss_utils.py:68-96 (batch_viewpoint_params_to_matrix) in fp32 numpy: x axis = normalised approach
    vector, y axis = (-x1, x0, 0) normalised ((0,1,0) when that is zero), z = x cross y, then an in-plane rotation
    by `angle`.  The reference calls torch.cross WITHOUT `dim`, which picks the FIRST dimension of size 3: with
    exactly three grasps the cross product runs down the rows instead of along each row (loss_utils.py:89) --
    restated as is."""
    f = np.float32
    ax = np.asarray(towards, f).reshape(-1, 3)
    n = ax.shape[0]
    ay = np.stack([-ax[:, 1], ax[:, 0], np.zeros(n, f)], axis=-1)
    ny = np.sqrt((ay * ay).sum(-1, dtype=f), dtype=f)
    ay[ny == 0, 1] = 1
    ax = ax / np.sqrt((ax * ax).sum(-1, dtype=f), dtype=f)[:, None]
    ay = ay / np.sqrt((ay * ay).sum(-1, dtype=f), dtype=f)[:, None]
    az = np.cross(ax, ay, axis=0 if n == 3 else -1).astype(f)
    s, c = np.sin(np.asarray(angle, f), dtype=f), np.cos(np.asarray(angle, f), dtype=f)
    R = np.empty((n, 3, 3), f)
    R[:, :, 0] = ax
    R[:, :, 1] = ay * c[:, None] + az * s[:, None]
    R[:, :, 2] = az * c[:, None] - ay * s[:, None]
    return R


def pred_decode(objectness_score, grasp_score_pred, grasp_angle_cls_pred, grasp_width_pred, grasp_tolerance_pred,
                fp2_xyz, grasp_top_view_xyz):
    """models/graspnet.py:77-134 for ONE scene: head outputs -> (M,17) fp32 grasp rows
    [score, width, 0.02, depth, R(9), centre(3), -1], seeds whose objectness argmax is 1, in seed order.
    Shapes: objectness (2,Ns); score / angle-class score / width / tolerance (A,Ns,D); fp2_xyz, top view (Ns,3).
    Everything but the rotation is a selection of inputs or a single-rounded fp32 expression (exact); the rotation
    goes through norms, sin/cos and a 3x3 product (compared to 1e-6)."""
    f = np.float32
    score = np.asarray(grasp_score_pred, f)
    cls = np.asarray(grasp_angle_cls_pred, f)
    width = np.clip(f(1.2) * np.asarray(grasp_width_pred, f), f(0), f(GRASP_MAX_WIDTH))      # :84-85
    tol = np.asarray(grasp_tolerance_pred, f)
    a_cls = np.argmax(cls, axis=0)                                                           # (Ns,D) first maximum
    angle = (a_cls.astype(f) / f(12)) * f(np.pi)                                             # :92
    pick = lambda t: np.take_along_axis(t, a_cls[None], axis=0)[0]
    score, width, tol = pick(score), pick(width), pick(tol)                                  # (Ns,D)
    d_cls = np.argmax(score, axis=1)[:, None]                                                # :101
    depth = (d_cls.astype(f) + f(1)) * f(0.01)
    take = lambda t: np.take_along_axis(t, d_cls, axis=1)
    score, angle, width, tol = take(score), take(angle), take(width), take(tol)              # (Ns,1)
    obj = np.asarray(objectness_score, f)
    mask = np.argmax(obj, axis=0) == 1                                                       # :110-111
    score, width, depth, angle, tol = score[mask], width[mask], depth[mask], angle[mask], tol[mask]
    centre = np.asarray(fp2_xyz, f)[mask]
    towards = -np.asarray(grasp_top_view_xyz, f)[mask]
    score = (score * tol) / f(GRASP_MAX_TOLERANCE)                                           # :119
    R = viewpoint_params_to_matrix(towards, angle.reshape(-1)).reshape(-1, 9)
    m = score.shape[0]
    return np.concatenate([score, width, np.full((m, 1), 0.02, f), depth, R, centre, np.full((m, 1), -1, f)],
                          axis=1).astype(f)
