"""One training step of GraspNet as the reference's train.py:120-143 runs it -- forward (models/graspnet.py:70-75 with
label generation), get_loss, backward, Adam -- at BASELINE.json config 5's shape (batch 8 per GPU, 20000 points),
under DistributedDataParallel when torch.distributed is initialised (one process per GPU, NCCL: the gradient
all-reduce is the only collective of the whole build).

    step = TrainStep(device, batch=8)            # synthetic clouds + labels with the dataset's keys and shapes
    loss = step.run()                            # forward + loss + backward + optimizer step

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 graspnet-baseline_b200/train_step.py --steps 5

In training mode the modules run the composed path: this library's geometry ops with their backward kernels
(group_points_grad, three_interpolate_grad, gather_points_grad) + torch layers with BatchNorm on batch statistics;
the labels come from the fused kernels of csrc/labels.cu (label_generation.FUSED)."""
import contextlib
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

if __package__ in (None, ""):
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import b200grasp  # noqa: F401
    from b200grasp import graspnet, label_generation, loss as loss_mod, pointnet2_utils, synth
else:
    from . import graspnet, label_generation, loss as loss_mod, pointnet2_utils, synth


def device_labels(clouds, num_objects, num_grasp_points, seed, num_view=300, num_angle=12, num_depth=4):
    """synth.training_labels with the big arrays drawn ON the device (14 GB per GPU at the dataset's 1200 grasp
    points per object would take minutes through numpy): same keys, shapes and value ranges"""
    dev = clouds.device
    g = torch.Generator(device=dev).manual_seed(seed)
    B, N, _ = clouds.shape
    V, A, D = num_view, num_angle, num_depth
    out = {k: [] for k in ("object_poses_list", "grasp_points_list", "grasp_offsets_list", "grasp_labels_list",
                           "grasp_tolerance_list")}
    obj = torch.zeros((B, N), dtype=torch.int64, device=dev)
    for b in range(B):
        scene = {k: [] for k in out}
        centres = clouds[b][torch.randperm(N, generator=g, device=dev)[:num_objects]]
        obj[b] = (torch.cdist(clouds[b], centres).min(dim=1)[0] < 0.08).long()
        for o in range(num_objects):
            q, _ = torch.linalg.qr(torch.randn(3, 3, generator=g, device=dev))
            if torch.det(q) < 0:
                q[:, 0] = -q[:, 0]
            scene["object_poses_list"].append(torch.cat([q, centres[o][:, None]], dim=1).contiguous())
            scene["grasp_points_list"].append(torch.randn(num_grasp_points, 3, generator=g, device=dev) * 0.03)
            shape = (num_grasp_points, V, A, D)
            lab = torch.rand(shape, generator=g, device=dev) * 1.2
            r = torch.rand(shape, generator=g, device=dev)
            lab[r < 0.7] = 0.0
            lab[r > 0.995] = -1.0
            off = torch.empty(shape + (3,), device=dev)
            off[..., 0] = (torch.arange(A, device=dev) * (np.pi / A))[None, None, :, None]
            off[..., 1] = (torch.arange(1, D + 1, device=dev) * 0.01)[None, None, None, :]
            off[..., 2] = torch.rand(shape, generator=g, device=dev) * 0.12
            scene["grasp_labels_list"].append(lab)
            scene["grasp_offsets_list"].append(off)
            scene["grasp_tolerance_list"].append(torch.rand(shape, generator=g, device=dev) * 0.05)
        for k in out:
            out[k].append(scene[k])
    out["objectness_label"] = obj
    return out


class TrainStep:
    def __init__(self, device, batch=8, num_point=20000, num_objects=5, num_grasp_points=300, seed=0, lr=1e-3,
                 ext=None, fused_labels=True):
        """ext: a module with the ten functions of pointnet2._ext to answer the pointnet2 ops instead of this library
        (the reference's legacy CUDA extension: the baseline arm); fused_labels: csrc/labels.cu or the composed chain"""
        self.device = torch.device(device)
        self.ext, self.fused_labels = ext, fused_labels
        rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        net = graspnet.GraspNet(input_feature_dim=0, num_view=300, num_angle=12, num_depth=4, cylinder_radius=0.05,
                                hmin=-0.02, hmax_list=[0.01, 0.02, 0.03, 0.04], is_training=True)
        synth.fill_state_dict(net, 12)                               # every rank the same initial parameters
        self.net = net.to(self.device).train()
        self.model = self.net
        if self.world > 1:
            self.model = torch.nn.parallel.DistributedDataParallel(
                self.net, device_ids=[self.device.index], output_device=self.device.index,
                gradient_as_bucket_view=True, bucket_cap_mb=8)       # ~4 MB of gradients: one bucket
        self.opt = torch.optim.Adam(self.net.parameters(), lr=lr)    # train.py:84
        clouds = np.stack([synth.tabletop_cloud(num_point, 1000 * (rank + 1) + seed + i) for i in range(batch)])
        self.clouds = torch.from_numpy(clouds).to(self.device)
        self.labels = device_labels(self.clouds, num_objects, num_grasp_points, seed=seed + 17 * rank)
        self.param_bytes = sum(p.numel() * 4 for p in self.net.parameters())

    def forward_loss(self):
        ep = dict(self.labels)
        ep["point_clouds"] = self.clouds
        old_ext, old_fused = pointnet2_utils._ext, label_generation.FUSED
        if self.ext is not None:
            pointnet2_utils._ext = self.ext
        label_generation.FUSED = self.fused_labels
        try:
            ep = self.model(ep)
        finally:
            pointnet2_utils._ext, label_generation.FUSED = old_ext, old_fused
        return loss_mod.get_loss(ep)

    def run(self, sync=True):
        """forward + loss + backward + Adam; sync=False skips the gradient all-reduce (DDP no_sync) -- for timing it"""
        self.opt.zero_grad(set_to_none=True)
        ctx = self.model.no_sync() if (not sync and self.world > 1) else contextlib.nullcontext()
        old_ext = pointnet2_utils._ext
        if self.ext is not None:
            pointnet2_utils._ext = self.ext                           # the backward calls the ops too
        try:
            with ctx:
                loss, _ = self.forward_loss()
                loss.backward()
        finally:
            pointnet2_utils._ext = old_ext
        self.opt.step()
        return loss.detach()

    def time(self, steps, warmup, sync=True):
        """-> milliseconds per step: CUDA events, barrier + synchronize on both sides, MAX over ranks"""
        for _ in range(warmup):
            self.run(sync)
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize(self.device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            self.run(sync)
        e1.record()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize(self.device)
        ms = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=self.device)
        if self.world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def allreduce_ms(self, reps=20):
        """the gradient all-reduce alone: one flat fp32 tensor of the parameters' size over NCCL"""
        if self.world == 1:
            return 0.0
        flat = torch.zeros(self.param_bytes // 4, dtype=torch.float32, device=self.device)
        for _ in range(3):
            dist.all_reduce(flat)
        torch.cuda.synchronize(self.device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            dist.all_reduce(flat)
        e1.record()
        torch.cuda.synchronize(self.device)
        return e0.elapsed_time(e1) / reps


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--grasp-points", type=int, default=300)
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    t0 = time.time()
    step = TrainStep(dev, batch=args.batch, num_grasp_points=args.grasp_points)
    ms = step.time(args.steps, args.warmup)
    ms_nosync = step.time(args.steps, 1, sync=False) if world > 1 else ms
    ar = step.allreduce_ms()
    if (dist.get_rank() if world > 1 else 0) == 0:
        print({"ms_per_step": round(ms, 2), "ms_per_step_without_allreduce": round(ms_nosync, 2), "allreduce_alone_ms": round(ar, 3),
               "scenes_per_s": round(world * args.batch / (ms * 1e-3), 1), "world": world, "batch_per_gpu": args.batch,
               "gradient_bytes": step.param_bytes, "setup_s": round(time.time() - t0, 1)})
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
