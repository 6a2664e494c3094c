#!/usr/bin/env python
"""One grid-walk ball query + cylinder query at batch B (argv[1], default 8) -- the command ncu captures."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
ops = b200grasp.ops
b = int(sys.argv[1]) if len(sys.argv) > 1 else 8
x = torch.from_numpy(np.stack([b200grasp.synth.tabletop_cloud(20000, 5 + i) for i in range(b)])).cuda()
idx = ops.furthest_point_sampling(x, 2048)
c = torch.gather(x, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
seeds = c[:, :1024].contiguous()
rot = torch.from_numpy(np.stack([b200grasp.synth.seed_rotations(1024, 9 + i) for i in range(b)])).reshape(b, 1024, 9).cuda()
for _ in range(2):
    ops.ball_query(c, x, 0.04, 64)
    ops.cylinder_query_multi(seeds, x, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64)
torch.cuda.synchronize()
print("ok")
