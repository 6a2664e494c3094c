"""One launch each of the HBM-bound kernels at the sizes where an HBM roofline is meaningful, for ncu --set full:
batch-64 group_points (SA2/SA3/SA4 features, SA1 / CloudCrop xyz), the cfg4 collision launch (4096 grasps x 200k points)
and the label-generation streaming pass (8 scenes).  Prints the algorithmic bytes of each launch."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200grasp
from b200grasp import ops, synth, label_generation as lg

dev = "cuda"
B = 64
g = torch.Generator(device=dev).manual_seed(0)
alg = {}
for name, c, n, m, ns in (("group_feat_sa2", 128, 2048, 1024, 32), ("group_feat_sa3", 256, 1024, 512, 16),
                          ("group_feat_sa4", 256, 512, 256, 16), ("group_xyz_sa1", 3, 20000, 2048, 64),
                          ("group_xyz_crop", 3, 20000, 4096, 64)):
    feats = torch.randn(B, c, n, device=dev, generator=g)
    idx = torch.randint(0, n, (B, m, ns), device=dev, generator=g, dtype=torch.int32)
    for _ in range(2):
        out = ops.group_points(feats, idx)
    alg[name] = B * (c * n * 4 + m * ns * 4 + c * m * ns * 4)
    del feats, idx, out
cloud64 = synth.tabletop_cloud(200000, 9, extent=0.6, dtype=np.float64)
gh = synth.synthetic_grasps(cloud64, 4096, 2)
det = object.__new__(b200grasp.ModelFreeCollisionDetector)
det.finger_width, det.finger_length = 0.01, 0.06
thr = torch.from_numpy(np.ascontiguousarray(det._thresholds(b200grasp.GraspGroup(gh), 0.05))).to(dev)
T = torch.from_numpy(np.ascontiguousarray(gh[:, 13:16], dtype=np.float64)).to(dev)
R = torch.from_numpy(np.ascontiguousarray(gh[:, 4:13], dtype=np.float64)).to(dev)
scene = torch.from_numpy(cloud64).to(dev)
for _ in range(2):
    ops.collision_counts(scene, T, R, thr)
alg["collision_cfg4"] = 200000 * 24 + 4096 * 22 * 8 + 4096 * 24
from b200grasp.train_step import device_labels
clouds = torch.from_numpy(np.stack([synth.tabletop_cloud(20000, 50 + i) for i in range(8)])).to(dev)
ep = dict(device_labels(clouds, 5, 300, seed=1))
inds = torch.stack([torch.randperm(20000, device=dev)[:1024] for _ in range(8)])
ep["fp2_xyz"] = torch.gather(clouds, 1, inds[..., None].expand(-1, -1, 3)).contiguous()
for _ in range(2):
    e = lg.process_grasp_labels(dict(ep), fused=True)
alg["label_view_stats"] = 8 * 1024 * 300 * (48 * 16 + 4)
torch.cuda.synchronize()
for k, v in alg.items():
    print("alg_bytes %-18s %d" % (k, v))
