"""CPU oracle for the fused group -> shared MLP -> max-pool op.  TEST INFRASTRUCTURE ONLY.

Restates, for eval mode,
  QueryAndGroup.forward / CylinderQueryAndGroup.forward with features=None, use_xyz=True
      (reference pointnet2/pointnet2_utils.py:345-350 and 523-530: group xyz, subtract the centre, optional
       division by the radius, optional row-vector x rotation matmul)
  -> pt_utils.SharedMLP (pointnet2/pytorch_utils.py:11-36, 60-116: Conv2d 1x1 without bias, BatchNorm2d, ReLU per layer)
  -> F.max_pool2d(kernel_size=[1, nsample]) (pointnet2/pointnet2_modules.py:246-249, models/modules.py:138-140)
  -> CloudCrop's view(B, -1, num_seed, num_depth) (models/modules.py:141).
Arithmetic is float64 on the float32 inputs with BatchNorm folded into (w, b) in float32 (fold_bn): the "exact"
value both the reference's fp32 evaluation and the CUDA kernel approximate.  Pinned by tests/golden/mlp_*.npz,
produced by running the reference's own SharedMLP (tests/golden/make_golden.py)."""
import numpy as np


def fold_bn(w, gamma, beta, mean, var, eps=1e-5):
    """conv weight (Cout,Cin) + eval-mode BatchNorm2d parameters -> (w', b') with y = w' x + b'"""
    scale = (gamma.astype(np.float32) / np.sqrt(var.astype(np.float32) + np.float32(eps))).astype(np.float32)
    return (w.astype(np.float32) * scale[:, None]).astype(np.float32), \
        (beta.astype(np.float32) - mean.astype(np.float32) * scale).astype(np.float32)


def grouped_xyz(xyz, new_xyz, idx, radius=0.0, rot=None):
    """xyz (B,N,3), new_xyz (B,M,3), idx (B,M,ns) -> (B,M,ns,3) float32, the reference's op order"""
    b = xyz.shape[0]
    g = np.stack([xyz[i][idx[i]] for i in range(b)]).astype(np.float32)          # (B,M,ns,3)
    g = g - new_xyz[:, :, None, :].astype(np.float32)
    if radius and radius > 0:
        g = (g / np.float32(radius)).astype(np.float32)
    if rot is not None:
        r = rot.reshape(b, -1, 3, 3).astype(np.float64)
        g = np.einsum("bmsj,bmjk->bmsk", g.astype(np.float64), r).astype(np.float32)
    return g


def group_mlp_maxpool(xyz, new_xyz, idx, layers, radius=0.0, rot=None, features=None):
    """idx (B,M,ns) or (G,B,M,ns); layers = [(w (Cout,Cin), b (Cout)), ...] already folded; features (B,C,N) or
    None: the first layer then sees torch.cat([grouped_xyz, grouped_features], dim=1) (pointnet2_utils.py:352-358)
    -> (B,C,M) or (B,C,M,G) float64"""
    squeeze = idx.ndim == 3
    lists = idx[None] if squeeze else idx
    b = xyz.shape[0]
    outs = []
    for g in range(lists.shape[0]):
        x = grouped_xyz(xyz, new_xyz, lists[g], radius, rot).astype(np.float64)   # (B,M,ns,3)
        if features is not None:
            gf = np.stack([features[i].T[lists[g][i]] for i in range(b)]).astype(np.float64)   # (B,M,ns,C)
            x = np.concatenate([x, gf], axis=-1)
        for w, bias in layers:
            x = np.maximum(x @ w.astype(np.float64).T + bias.astype(np.float64), 0.0)
        outs.append(x.max(axis=2).transpose(0, 2, 1))                           # (B,C,M)
    out = np.stack(outs, axis=-1)
    return out[..., 0] if squeeze else out
