#!/usr/bin/env python
"""Time the SA1 ball query and the CloudCrop cylinder query, grid walk vs linear scan, batch 1 and 64."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
ops = b200grasp.ops

def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3

for b in (1, 64):
    x = torch.from_numpy(np.stack([b200grasp.synth.tabletop_cloud(20000, 5 + i) for i in range(b)])).cuda()
    idx = ops.furthest_point_sampling(x, 2048)
    c = torch.gather(x, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
    seeds = c[:, :1024].contiguous()
    rot = torch.from_numpy(np.stack([b200grasp.synth.seed_rotations(1024, 9 + i) for i in range(b)])).reshape(b, 1024, 9).cuda()
    for name, thr, nt in (("grid64", 4096, 64), ("grid128", 4096, 128), ("grid256", 4096, 256), ("linear", 1 << 30, 0)):
        ops.GRID_QUERY_MIN_N = thr
        b200grasp._abi.lib.b200_query_grid_set_threads(nt)
        ub = t(lambda: ops.ball_query(c, x, 0.04, 64))
        uc = t(lambda: ops.cylinder_query_multi(seeds, x, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64))
        print("batch %2d %-7s: ball_sa1 %8.1f us   cylinder_multi %8.1f us" % (b, name, ub, uc), flush=True)

print("shared grid (one binning for both queries): cell edge sweep")
lib = b200grasp._abi.lib
for b in (1, 64):
    x = torch.from_numpy(np.stack([b200grasp.synth.tabletop_cloud(20000, 5 + i) for i in range(b)])).cuda()
    idx = ops.furthest_point_sampling(x, 2048)
    c = torch.gather(x, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
    seeds = c[:, :1024].contiguous()
    rot = torch.from_numpy(np.stack([b200grasp.synth.seed_rotations(1024, 9 + i) for i in range(b)])).reshape(b, 1024, 9).cuda()
    for cell in (0.016, 0.0202, 0.026, 0.0323, 0.045):
        for nt in ((256, 128) if b == 1 else (128, 64)):
            lib.b200_query_grid_set_threads(nt)
            g = ops.query_grid_build(x, cell)
            ug = t(lambda: ops.query_grid_build(x, cell))
            ub = t(lambda: ops.ball_query(c, x, 0.04, 64, grid=g))
            uc = t(lambda: ops.cylinder_query_multi(seeds, x, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64, grid=g))
            print("batch %2d cell %.4f nt %3d: build %7.1f us  ball_sa1 %7.1f us  cylinder_multi %7.1f us" % (b, cell, nt, ug, ub, uc), flush=True)
lib.b200_query_grid_set_threads(0)
