"""One CloudCrop-sized (or SA1-sized: argv[1] == 'sa1') launch of the fused group -> MLP -> max-pool kernel, for ncu."""
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200grasp
from b200grasp import ops, fused_modules as fm

which = sys.argv[1] if len(sys.argv) > 1 else "crop"
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 3
xyz = torch.from_numpy(b200grasp.synth.tabletop_cloud(20000, 5)[None]).cuda()
torch.manual_seed(0)
inds = ops.furthest_point_sampling(xyz, 2048)
new_xyz = ops.gather_xyz(xyz, inds, transposed=False)
with torch.no_grad():
    if which == "sa1":
        mod = fm.PointnetSAModuleVotes(npoint=2048, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], normalize_xyz=True).cuda().eval()
        idx = ops.ball_query(new_xyz, xyz, 0.04, 64)
        layers = mod._fused.get()
        for _ in range(3):
            out = ops.group_mlp_maxpool(xyz, new_xyz, idx, layers, radius=0.04, passes=passes)
    else:
        mod = fm.CloudCrop(64, 3).cuda().eval()
        seeds = new_xyz[:, :1024].contiguous()
        rot = torch.from_numpy(b200grasp.synth.seed_rotations(1024, 9).reshape(1, 1024, 9)).cuda()
        idx = ops.cylinder_query_multi(seeds, xyz, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64)
        layers = mod._fused.get()
        for _ in range(3):
            out = ops.group_mlp_maxpool(xyz, seeds, idx, layers, rot=rot, passes=passes)
torch.cuda.synchronize()
print(which, passes, tuple(out.shape), float(out.abs().max()))
