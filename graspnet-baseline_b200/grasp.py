"""Minimal GraspGroup: the accessors the hot path reads, plus nms on the GPU.

graspnetAPI is an EMPTY, un-vendored submodule in the reference;
this mirrors the upstream container's surface as the reference uses it: the
17-column row layout fixed by models/graspnet.py:131-133, the accessors read by
utils/collision_detector.py:70-74, gg[mask] (test.py:96), len(), sort_by_score()
(demo_pointcloud.py:250), nms() (demo_pointcloud.py:249), save_npy (test.py:103).
"""
import math

import numpy as np
import torch

from . import ops as _ext

GRASP_ARRAY_LEN = 17


class GraspGroup:
    def __init__(self, *args):
        if len(args) == 0:
            self.grasp_group_array = np.zeros((0, GRASP_ARRAY_LEN), dtype=np.float64)
        elif len(args) == 1 and isinstance(args[0], np.ndarray):
            self.grasp_group_array = args[0]
        elif len(args) == 1 and isinstance(args[0], str):
            self.grasp_group_array = np.load(args[0])
        else:
            raise ValueError("args must be nothing, numpy array or string.")

    def __len__(self):
        return len(self.grasp_group_array)

    def __getitem__(self, index):
        if isinstance(index, (int, np.integer)):
            return self.grasp_group_array[index]
        if isinstance(index, (slice, list, np.ndarray)):
            return GraspGroup(self.grasp_group_array[index])
        raise TypeError("unknown type %s for GraspGroup index" % type(index))

    scores = property(lambda s: s.grasp_group_array[:, 0])
    widths = property(lambda s: s.grasp_group_array[:, 1])
    heights = property(lambda s: s.grasp_group_array[:, 2])
    depths = property(lambda s: s.grasp_group_array[:, 3])
    rotation_matrices = property(lambda s: s.grasp_group_array[:, 4:13].reshape((-1, 3, 3)))
    translations = property(lambda s: s.grasp_group_array[:, 13:16])
    object_ids = property(lambda s: s.grasp_group_array[:, 16])

    def sort_by_score(self, reverse=False):
        """In place; descending unless reverse (stable, so equal scores keep their order)."""
        score = self.grasp_group_array[:, 0]
        index = np.argsort(score if reverse else -score, kind="stable")
        self.grasp_group_array = self.grasp_group_array[index]
        return self

    def save_npy(self, npy_file_path):
        np.save(npy_file_path, self.grasp_group_array)

    def nms(self, translation_thresh=0.03, rotation_thresh=30.0 / 180.0 * math.pi, device=None):
        """Greedy score-ordered suppression; returns a NEW GraspGroup (rows in score order)."""
        if len(self) == 0:
            return GraspGroup(self.grasp_group_array.copy())
        dev = torch.device(device if device is not None else "cuda")
        g = torch.from_numpy(np.ascontiguousarray(self.grasp_group_array, dtype=np.float32)).to(dev)
        keep, nkeep = _ext.grasp_nms(g, translation_thresh, math.cos(rotation_thresh))
        both = torch.cat([nkeep, keep]).cpu().numpy()          # one D2H copy, one synchronisation
        rows = both[1: 1 + int(both[0])]
        return GraspGroup(self.grasp_group_array[rows])
