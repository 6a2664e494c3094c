"""ctypes binding of libb200grasp.so (include/b200grasp.h).

The CUDA library is the ONLY implementation: if it is missing this module raises
at import, there is no CPU or PyTorch fallback anywhere in the package.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200grasp.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "b200grasp: %s not found -- build it with `python graspnet-baseline_b200/build.py` "
        "(nvcc, sm_100a). There is no fallback path." % LIB_PATH)

lib = ctypes.CDLL(LIB_PATH)

_vp, _i, _f, _d, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_double, ctypes.c_size_t


class MlpLayer(ctypes.Structure):
    """b200_mlp_layer (include/b200grasp.h): one tensor-core layer of the fp16x3 kernels"""
    _fields_ = [("w_packed", _vp), ("bias", _vp), ("k_pad", _i), ("n_pad", _i), ("cout", _i), ("relu", _i),
                ("w_exp", _i), ("w_norm1", _f), ("b_max", _f), ("y", _vp), ("y_batch_stride", ctypes.c_longlong)]


_lp = ctypes.POINTER(MlpLayer)

# name -> argtypes, in the order of include/b200grasp.h
SIGNATURES = {
    "b200_furthest_point_sampling": [_vp, _i, _i, _i, _vp, _vp],
    "b200_gather_points": [_vp, _vp, _i, _i, _i, _i, _vp, _vp],
    "b200_gather_points_grad": [_vp, _vp, _i, _i, _i, _i, _vp, _vp],
    "b200_ball_query": [_vp, _vp, _i, _i, _i, _f, _i, _vp, _vp],
    "b200_cylinder_query": [_vp, _vp, _vp, _i, _i, _i, _f, _f, _f, _i, _vp, _vp],
    "b200_cylinder_query_multi": [_vp, _vp, _vp, _i, _i, _i, _f, _f, ctypes.POINTER(_f), _i, _i, _vp, _vp],
    "b200_group_points": [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp],
    "b200_group_points_grad": [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp],
    "b200_three_nn": [_vp, _vp, _i, _i, _i, _vp, _vp, _vp],
    "b200_three_interpolate": [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp],
    "b200_three_interpolate_grad": [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp],
    "b200_knn": [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp],
    "b200_collision_counts": [_vp, _i, _vp, _vp, _vp, _i, _vp, _vp],
    "b200_grasp_nms": [_vp, _i, _d, _d, _vp, _vp, _vp, _sz, _vp],
    "b200_label_object_table_set": [_vp, _i, _vp, _vp, _vp, _vp, _vp],
    "b200_label_view_stats": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _vp, _vp, _vp],
    "b200_label_view_scores": [_vp, _vp, ctypes.c_longlong, _vp, _vp, _vp],
    "b200_label_top_gather": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp],
    "b200_grasp_nms_static": [_vp, _i, _d, _d, _vp, _vp, _vp],
    "b200_grasp_post_chain": [_vp, _vp, _i, _vp, _vp, _i, _d, _d, _d, _d, _d, _vp, _vp, _vp, _vp, _vp, _sz, _vp],
    "b200_collision_counts_dn": [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp],
    "b200_voxel_down_sample": [_vp, _i, _d, _vp, _vp, _vp, _sz, _vp],
    "b200_voxel_down_sample_checked": [_vp, _i, _d, _vp, _vp, _vp, _vp, _sz, _vp],
    "b200_collision_counts_packed": [_vp, _vp, _i, _vp, _i, _vp, _vp],
    "b200_collision_mask_device": [_vp, _vp, _i, _vp, _vp, _i, _d, _d, _d, _vp, _vp, _sz, _vp],
    "b200_compact_rows": [_vp, _i, _vp, _vp, _i, _vp, _vp, _vp],
    "b200_grasp_nms_device": [_vp, _vp, _i, _d, _d, _vp, _vp, _vp, _sz, _vp],
    "b200_gather_rows": [_vp, _i, _vp, _vp, _i, _vp, _vp],
    "b200_fps_prefix_check": [_vp, _i, _i, _i, ctypes.c_longlong, _vp, _vp],
    "b200_furthest_point_sampling_spec": [_vp, _i, _i, _i, _vp, _i, _vp, _i, _vp, _i, _vp, _vp],
    "b200_group_mlp_maxpool": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp],
    "b200_group_feat_mlp_maxpool": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp],
    "b200_mlp_pack_weights": [_vp, _i, _i, _vp, _vp],
    "b200_gather_xyz": [_vp, _vp, _i, _i, _i, _vp, _vp, _vp],
    "b200_pred_decode": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp],
    "b200_pointwise_conv": [_vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _vp, _vp],
    "b200_view_select": [_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "b200_three_nn_weights": [_vp, _i, _vp, _vp, _vp],
    "b200_ball_query_grid": [_vp, _vp, _i, _i, _i, _f, _i, _vp, _vp, _sz, _vp],
    "b200_cylinder_query_multi_grid": [_vp, _vp, _vp, _i, _i, _i, _f, _f, ctypes.POINTER(_f), _i, _i, _vp, _vp, _sz, _vp],
    "b200_query_grid_build": [_vp, _i, _i, _f, _vp, _sz, _vp],
    "b200_ball_query_prebuilt": [_vp, _vp, _i, _i, _i, _f, _i, _vp, _vp, _sz, _vp],
    "b200_mlp_f16_pack": [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp],
    "b200_group_mlp_maxpool_f16": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _vp, _vp, _i, _f, _f, _lp, _vp, _vp, _vp, _vp],
    "b200_group_feat_mlp_maxpool_f16": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _lp, _vp, _vp, _vp, _vp],
    "b200_pointwise_stack_f16": [_vp, _i, _vp, _i, _i, _i, _lp, _i, _i, _vp],
    "b200_cylinder_query_multi_prebuilt": [_vp, _vp, _vp, _i, _i, _i, _f, _f, ctypes.POINTER(_f), _i, _i, _vp, _vp, _sz, _vp],
}
for _name, _args in SIGNATURES.items():
    _fn = getattr(lib, _name)
    _fn.argtypes = _args
    _fn.restype = _i
lib.b200_grasp_nms_workspace.argtypes = [_i]
lib.b200_grasp_nms_workspace.restype = _sz
lib.b200_voxel_down_sample_workspace.argtypes = [_i]
lib.b200_voxel_down_sample_workspace.restype = _sz
lib.b200_collision_mask_workspace.argtypes = [_i]
lib.b200_collision_mask_workspace.restype = _sz
lib.b200_grasp_post_chain_workspace.argtypes = [_i]
lib.b200_grasp_post_chain_workspace.restype = _sz
lib.b200_label_object_table_bytes.argtypes = [_i]
lib.b200_label_object_table_bytes.restype = _sz
lib.b200_mlp_f16_packed_bytes.argtypes = [_i, _i]
lib.b200_mlp_f16_packed_bytes.restype = _sz
lib.b200_query_grid_workspace.argtypes = [_i, _i]
lib.b200_query_grid_workspace.restype = _sz
lib.b200_error_string.argtypes = [_i]
lib.b200_error_string.restype = ctypes.c_char_p
lib.b200_version.restype = _i
lib.b200_launch_count.restype = ctypes.c_ulonglong
lib.b200_fps_set_cluster.argtypes = [_i]
lib.b200_fps_set_cluster.restype = None
lib.b200_fps_set_max_cluster.argtypes = [_i]
lib.b200_fps_set_max_cluster.restype = None
lib.b200_fps_set_threads.argtypes = [_i]
lib.b200_fps_set_threads.restype = None
lib.b200_query_grid_set_threads.argtypes = [_i]
lib.b200_query_grid_set_threads.restype = None
lib.b200_fps_set_multi.argtypes = [_i]
lib.b200_fps_set_multi.restype = None
lib.b200_grasp_nms_set_options.argtypes = [ctypes.c_uint]
lib.b200_grasp_nms_set_options.restype = None
lib.b200_voxel_set_options.argtypes = [ctypes.c_uint]
lib.b200_voxel_set_options.restype = None
NMS_TRANS_INCLUSIVE, NMS_ROT_INCLUSIVE, NMS_SAME_OBJECT, NMS_FP32, NMS_ACOS = 1, 2, 4, 8, 16
VOXEL_ORIGIN_AT_MIN = 1
VOXEL_MULTI_CTA = 2
lib.b200_mlp_f16_set_tiles_per_cta.argtypes = [_i]
lib.b200_mlp_f16_set_tiles_per_cta.restype = None
lib.b200_fps_set_speculation.argtypes = [_i]
lib.b200_fps_set_speculation.restype = None

EXPORTS = sorted(list(SIGNATURES) + ["b200_grasp_nms_workspace", "b200_voxel_down_sample_workspace", "b200_query_grid_workspace", "b200_collision_mask_workspace", "b200_grasp_post_chain_workspace", "b200_mlp_f16_packed_bytes", "b200_label_object_table_bytes", "b200_grasp_nms_set_options", "b200_voxel_set_options", "b200_mlp_f16_set_tiles_per_cta", "b200_error_string", "b200_version", "b200_launch_count"])


def check(code, what):
    """Non-zero status -> RuntimeError (the reference exit(-1)s instead, cuda_utils.h:35-44)."""
    if code != 0:
        raise RuntimeError("%s failed: %s (code %d)" % (what, lib.b200_error_string(code).decode(), code))

POST_CHAIN_MAX_M = 2048                # B200_POST_CHAIN_MAX_M (include/b200grasp.h)
