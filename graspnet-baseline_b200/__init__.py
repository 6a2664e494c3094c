"""b200grasp -- B200 (sm_100a) implementation of GraspNet-baseline's point-cloud
geometry hot path behind the reference's own call signatures.

    import b200grasp                      # repo-root alias of this directory
    from b200grasp import pointnet2_utils, knn_modules, collision_detector, grasp
    b200grasp.install_dropins()           # makes `import pointnet2_utils`, `from knn_modules import knn`,
                                          # `from collision_detector import ...`, `from graspnetAPI import GraspGroup`
                                          # resolve to this package, as the reference's sys.path-style imports expect

Importing this package loads lib/libb200grasp.so and fails loudly if it is absent.
"""
import os
import sys

from . import _abi  # noqa: F401  (raises ImportError if the CUDA library is missing)
from . import ops, pointnet2_utils, knn_modules, grasp, voxel, collision_detector, decode, postprocess, synth, sharding, fused_modules, graspnet, label_generation, loss, custom_ops  # noqa: F401
from .grasp import GraspGroup  # noqa: F401
from .collision_detector import ModelFreeCollisionDetector  # noqa: F401
from .decode import pred_decode, pred_decode_device  # noqa: F401
from .postprocess import grasp_postprocess, grasp_postprocess_device  # noqa: F401

__version__ = "0.1.0"
LIB_PATH = _abi.LIB_PATH


def install_dropins():
    """Register the reference's bare module names (pointnet2/pointnet2_modules.py:18-22,
    models/modules.py:14-18, utils/label_generation.py:12-14 import them that way)."""
    import types
    sys.modules["pointnet2_utils"] = pointnet2_utils
    sys.modules["knn_modules"] = knn_modules
    sys.modules["collision_detector"] = collision_detector
    sys.modules["label_generation"] = label_generation       # models/graspnet.py:20
    sys.modules["loss"] = loss                                # train.py:26
    api = types.ModuleType("graspnetAPI")
    api.GraspGroup = GraspGroup
    sys.modules["graspnetAPI"] = api
