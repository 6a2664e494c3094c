"""Voxel down-sample behind ModelFreeCollisionDetector.__init__ (the reference calls
open3d's PointCloud.voxel_down_sample, utils/collision_detector.py:32-35; open3d is
not available here and its arithmetic is restated: origin = min_bound - voxel/2,
key = floor((p - origin)/voxel), output = fp64 mean per occupied voxel accumulated
in point order, rows sorted by key -- PARITY UNPINNED).

Host-side numpy for now (constructor cost, outside the detect() hot loop);
a CUDA version is row f3 of SURVEY.md section 8(f).
"""
import numpy as np


def voxel_down_sample(points, voxel_size):
    p = np.asarray(points, dtype=np.float64).reshape(-1, 3)
    if p.shape[0] == 0:
        return p
    origin = p.min(axis=0) - voxel_size * 0.5
    key = np.floor((p - origin) / voxel_size).astype(np.int64)
    order = np.lexsort((key[:, 2], key[:, 1], key[:, 0]))
    ks = key[order]
    first = np.ones(len(ks), dtype=bool)
    first[1:] = np.any(ks[1:] != ks[:-1], axis=1)
    seg = np.cumsum(first) - 1
    nvox = int(seg[-1]) + 1
    out = np.zeros((nvox, 3), dtype=np.float64)
    np.add.at(out, seg, p[order])            # unbuffered, in order => sequential accumulation
    cnt = np.bincount(seg, minlength=nvox).astype(np.float64)
    return out / cnt[:, None]
