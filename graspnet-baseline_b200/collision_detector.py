"""Drop-in for the reference's utils/collision_detector.py (ModelFreeCollisionDetector).

Same constructor and detect() signature / return conventions
(collision_detector.py:28-35, 37-129); numpy in, numpy out, CUDA inside.
The M x N point-in-gripper tests run in libb200grasp.so (fp64, integer counts);
the per-grasp thresholds and the count -> IoU -> mask arithmetic are M-length
numpy expressions with the reference's own dtype rules, so results are
bit-identical to the reference (tests/golden/collision_*.npz).
"""
import numpy as np
import torch

from . import ops as _ext
from .voxel import check_voxel_input, voxel_down_sample_device


class ModelFreeCollisionDetector():
    def __init__(self, scene_points, voxel_size=0.005, device=None):
        self.finger_width = 0.01
        self.finger_length = 0.06
        self.voxel_size = voxel_size
        self.device = torch.device(device if device is not None else "cuda")
        # open3d voxel_down_sample in the reference (:32-35); here on the device, no host round trip:
        # detect() reads the row count from device memory
        self._scene_dev, self._nvox_dev, self._nbad_dev = voxel_down_sample_device(scene_points, voxel_size, self.device,
                                                                                   return_bad=True)
        self._scene_host = None

    @property
    def scene_points(self):
        """(Nv,3) fp64 numpy, as the reference attribute (copied back on first use)."""
        if self._scene_host is None:
            n = self._scene_dev.shape[0] if self._nvox_dev is None else int(self._nvox_dev.item())
            self._scene_host = self._scene_dev[:n].cpu().numpy()
        return self._scene_host

    @scene_points.setter
    def scene_points(self, value):
        self._scene_host = np.asarray(value, dtype=np.float64)
        self._scene_dev = torch.from_numpy(np.ascontiguousarray(self._scene_host)).to(self.device)
        self._nvox_dev = self._nbad_dev = None

    def _thresholds(self, grasp_group, approach_dist):
        h = grasp_group.heights
        d = grasp_group.depths
        w = grasp_group.widths
        fl, fw = self.finger_length, self.finger_width
        cols = [-h / 2, h / 2, d - fl, d, -(w / 2 + fw), -w / 2, (w / 2 + fw), w / 2,
                d - fl - fw, d - fl - fw - approach_dist]
        return np.stack([np.asarray(c, dtype=np.float64) for c in cols], axis=1)

    def detect(self, grasp_group, approach_dist=0.03, collision_thresh=0.05, return_empty_grasp=False,
               empty_thresh=0.01, return_ious=False):
        approach_dist = max(approach_dist, self.finger_width)
        m = len(grasp_group)
        # one packed host buffer (T | R | thr) -> one H2D copy; one D2H copy of the integer counts
        packed = np.empty((m, 22), dtype=np.float64)
        packed[:, 0:3] = grasp_group.translations
        packed[:, 3:12] = grasp_group.rotation_matrices.reshape(-1, 9)
        packed[:, 12:22] = self._thresholds(grasp_group, approach_dist)
        d = torch.from_numpy(packed).to(self.device, non_blocking=True)
        counts = _ext.collision_counts_packed(self._scene_dev, d, self._nvox_dev).cpu().numpy().astype(np.int64)
        if self._nbad_dev is not None:                   # (the constructor does not synchronise: checked at first use)
            check_voxel_input(self._nbad_dev)
            self._nbad_dev = None
        heights = grasp_group.heights[:, np.newaxis]
        widths = grasp_group.widths[:, np.newaxis]
        vox3 = self.voxel_size ** 3
        left_right_volume = (heights * self.finger_length * self.finger_width / vox3).reshape(-1)
        bottom_volume = (heights * (widths + 2 * self.finger_width) * self.finger_width / vox3).reshape(-1)
        shifting_volume = (heights * (widths + 2 * self.finger_width) * approach_dist / vox3).reshape(-1)
        volume = left_right_volume * 2 + bottom_volume + shifting_volume
        global_iou = counts[:, 0] / (volume + 1e-6)
        collision_mask = (global_iou > collision_thresh)
        if not (return_empty_grasp or return_ious):
            return collision_mask
        ret_value = [collision_mask, ]
        if return_empty_grasp:
            inner_volume = (heights * self.finger_length * widths / vox3).reshape(-1)
            ret_value.append(counts[:, 5] / inner_volume < empty_thresh)
        if return_ious:
            ret_value.append([global_iou, counts[:, 1] / (left_right_volume + 1e-6),
                              counts[:, 2] / (left_right_volume + 1e-6), counts[:, 3] / (bottom_volume + 1e-6),
                              counts[:, 4] / (shifting_volume + 1e-6)])
        return ret_value
