"""Drop-in for the reference's knn/knn_modules.py (knn(ref, query, k=1), :9-17)."""
import torch

from . import ops as _ext


def knn(ref, query, k=1):
    """k nearest reference points for each query point.
    ref (B,D,Nr), query (B,D,Nq) -> (B,k,Nq) int64, 1-BASED (callers subtract 1,
    utils/label_generation.py:60,86)."""
    device = ref.device
    ref = ref.float().to(device).contiguous()
    query = query.float().to(device).contiguous()
    inds = torch.empty(query.shape[0], k, query.shape[2], dtype=torch.long, device=device)
    _ext.knn(ref, query, inds)
    return inds
