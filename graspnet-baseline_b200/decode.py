"""Drop-in for pred_decode (reference models/graspnet.py:77-134): network head outputs -> grasp rows.

`pred_decode(end_points)` keeps the reference's signature and return value (a list with one (M_i,17) tensor per
scene, on the inputs' device); `pred_decode_device(end_points)` returns the padded rows and the per-scene counts
without synchronising, for callers that keep post-processing on the device (collision + NMS).
"""
import torch

from . import ops as _ext

_KEYS = ("objectness_score", "grasp_score_pred", "grasp_angle_cls_pred", "grasp_width_pred", "grasp_tolerance_pred",
         "fp2_xyz", "grasp_top_view_xyz")


def _inputs(end_points):
    # the reference casts with .float() (graspnet.py:82-85) and indexes per scene; contiguity is free for head outputs
    return [end_points[k].float().contiguous() for k in _KEYS]


def pred_decode_device(end_points):
    """-> (rows (B,Ns,17) f32, nrows (B,) i32) on the device; rows[i, :nrows[i]] are scene i's grasps."""
    return _ext.pred_decode(*_inputs(end_points))


def pred_decode(end_points):
    rows, nrows = pred_decode_device(end_points)
    counts = nrows.cpu().tolist()                      # the reference's boolean-mask indexing synchronises too
    return [rows[i, :n] for i, n in enumerate(counts)]
