"""A few CloudCrop-sized (or sa1 / sa2 / heads: argv[1]) launches of the fp16x3 kernels (csrc/mlp_f16.cu), for ncu."""
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200grasp
from b200grasp import ops, fused_modules as fm

which = sys.argv[1] if len(sys.argv) > 1 else "crop"
xyz = torch.from_numpy(b200grasp.synth.tabletop_cloud(20000, 5)[None]).cuda()
torch.manual_seed(0)
inds = ops.furthest_point_sampling(xyz, 2048)
new_xyz = ops.gather_xyz(xyz, inds, transposed=False)
with torch.no_grad():
    if which == "crop":
        mod = fm.CloudCrop(64, 3).cuda().eval()
        seeds = new_xyz[:, :1024].contiguous()
        rot = torch.from_numpy(b200grasp.synth.seed_rotations(1024, 9).reshape(1, 1024, 9)).cuda()
        idx = ops.cylinder_query_multi(seeds, xyz, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64)
        first, l16 = mod._fused.get_f16()
        run = lambda: ops.group_mlp_maxpool_f16(xyz, seeds, idx, first, l16, rot=rot)
    elif which == "sa1":
        mod = fm.PointnetSAModuleVotes(npoint=2048, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], normalize_xyz=True).cuda().eval()
        idx = ops.ball_query(new_xyz, xyz, 0.04, 64)
        first, l16 = mod._fused.get_f16()
        run = lambda: ops.group_mlp_maxpool_f16(xyz, new_xyz, idx, first, l16, radius=0.04)
    elif which == "sa2":
        mod = fm.PointnetSAModuleVotes(npoint=1024, radius=0.1, nsample=32, mlp=[128, 128, 128, 256], normalize_xyz=True).cuda().eval()
        xyz2 = new_xyz[:, :1024].contiguous()
        idx = ops.ball_query(xyz2, new_xyz, 0.1, 32)
        feat_t = torch.randn(1, 2048, 128, device="cuda")
        l16 = mod._fused.get_f16()[1]
        run = lambda: ops.group_feat_mlp_maxpool_f16(new_xyz, xyz2, feat_t, idx, l16, radius=0.1)
    else:
        g = torch.Generator(device="cuda").manual_seed(1)
        dims = [256, 128, 128, 36]
        ws = [(torch.randn(co, ci, device="cuda", generator=g) * (2.0 / ci) ** 0.5, torch.randn(co, device="cuda", generator=g) * 0.1)
              for ci, co in zip(dims[:-1], dims[1:])]
        x = torch.relu(torch.randn(1, 256, 4096, device="cuda", generator=g))
        la = ops.f16_stack(ws, relu_last=False)
        lb = ops.f16_stack(ws, relu_last=False)
        run = lambda: ops.pointwise_stack_f16(x, la, layers2=lb)
    for _ in range(4):
        out = run()
torch.cuda.synchronize()
print(which, "ok")
