"""Scene-level sharding across GPUs: scenes are independent, so rank r of G takes
scenes r, r+G, r+2G, ... and nothing is exchanged on the data path.  The only collectives are the two scalar
reductions bench.py needs to report a whole-job number: total scenes (SUM) and the slowest rank's time (MAX)."""
import torch
import torch.distributed as dist


def scene_shard(num_scenes, rank, world):
    """Scene ids processed by `rank` (round-robin, like `scene s -> GPU s mod G`)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return list(range(rank, num_scenes, world))


def whole_job_throughput(local_scenes, local_ms, device="cpu"):
    """(total scenes, max-over-ranks milliseconds, scenes/s) over the default process group
    (or the local values when torch.distributed is not initialised)."""
    t = torch.tensor([float(local_ms)], dtype=torch.float64, device=device)
    n = torch.tensor([float(local_scenes)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(n, op=dist.ReduceOp.SUM)
    ms, total = float(t.item()), float(n.item())
    return total, ms, (total / (ms * 1e-3) if ms > 0 else 0.0)
