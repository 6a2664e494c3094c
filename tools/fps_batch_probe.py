#!/usr/bin/env python
"""FPS on a batch of FPS-ordered prefixes (SA2 shape at batch 64): time with/without speculation."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
lib = b200grasp._abi.lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
x = torch.from_numpy(np.stack([b200grasp.synth.tabletop_cloud(20000, 100 + i) for i in range(B)])).cuda()
idx = b200grasp.ops.furthest_point_sampling(x, 2048)
p = torch.gather(x, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
for spec in (1, 0):
    lib.b200_fps_set_speculation(spec)
    for _ in range(2):
        out = b200grasp.ops.furthest_point_sampling(p, 1024)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); out = b200grasp.ops.furthest_point_sampling(p, 1024); e1.record(); torch.cuda.synchronize()
    ident = (out.cpu().numpy() == np.arange(1024)[None]).all(axis=1)
    print("B=%d spec=%d: %.1f us, identity in %d/%d clouds" % (B, spec, e0.elapsed_time(e1) * 1e3, ident.sum(), B))
