"""N>1 training plumbing on CPU: a world_size-2 gloo run of DistributedDataParallel over the part of the training
step that is plain torch -- the two grasp heads (OperationNet / ToleranceNet in training mode) and the grasp loss --
checking that every rank ends up with the gradient of the two ranks' mean loss (what the NCCL all-reduce of
train_step.TrainStep does on the GPUs; the geometry ops of the rest of the network are CUDA-only)."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def rank_batch(b200, rank, ns=32):
    rng = np.random.default_rng(100 + rank)
    f = np.float32
    ep = {"objectness_label": torch.from_numpy((rng.random((1, 400)) < 0.6).astype(np.int64)),
          "fp2_inds": torch.from_numpy(rng.choice(400, ns, replace=False)[None].astype(np.int32)),
          "batch_grasp_label": torch.from_numpy(rng.uniform(0, 2, (1, ns, 12, 4)).astype(f)),
          "batch_grasp_offset": torch.from_numpy(rng.uniform(0, 0.1, (1, ns, 12, 4, 3)).astype(f)),
          "batch_grasp_tolerance": torch.from_numpy(rng.uniform(0, 0.05, (1, ns, 12, 4)).astype(f)),
          "grasp_top_view_inds": torch.zeros((1, ns), dtype=torch.int64)}
    return ep, torch.from_numpy(rng.standard_normal((1, 256, ns, 4)).astype(f))


class Heads(torch.nn.Module):
    def __init__(self, b200):
        super().__init__()
        self.operation = b200.graspnet.OperationNet(12, 4)
        self.tolerance = b200.graspnet.ToleranceNet(12, 4)

    def forward(self, vp_features, end_points):
        return self.tolerance(vp_features, self.operation(vp_features, end_points))


def grads_of(b200, model, batches):
    model.zero_grad()
    total = 0
    for ep, x in batches:
        loss, _ = b200.loss.compute_grasp_loss(model(x, dict(ep)))
        total = total + loss / len(batches)
    total.backward()
    return float(total)


def worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import b200grasp as b200
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    heads = Heads(b200).train()
    b200.synth.fill_state_dict(heads, 3)
    ddp = torch.nn.parallel.DistributedDataParallel(heads)
    grads_of(b200, ddp, [rank_batch(b200, rank)])
    q.put((rank, {k: p.grad.clone().numpy() for k, p in heads.named_parameters()}))
    dist.destroy_process_group()


def test_two_rank_ddp_gradients_are_the_mean_of_the_ranks():
    import b200grasp as b200
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    torch.manual_seed(0)
    heads = Heads(b200).train()
    b200.synth.fill_state_dict(heads, 3)
    # (BatchNorm uses each rank's own batch statistics under DDP: the single-process statement runs the ranks' batches
    # one after the other, not concatenated)
    grads_of(b200, heads, [rank_batch(b200, 0), rank_batch(b200, 1)])
    for k, p in heads.named_parameters():
        want = p.grad.numpy()
        scale = max(float(np.abs(want).max()), 1e-8)
        for r in (0, 1):
            assert np.abs(res[r][k] - want).max() <= 1e-5 * scale, (k, r)
        assert (res[0][k] == res[1][k]).all()                 # both ranks hold the same reduced gradient
