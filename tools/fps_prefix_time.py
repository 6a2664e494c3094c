#!/usr/bin/env python
"""Time FPS on FPS-ordered prefixes (SA2..SA4 shapes) with and without the speculative check."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
lib = b200grasp._abi.lib

def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3

x = torch.from_numpy(b200grasp.synth.tabletop_cloud(20000, 5)[None]).cuda()
for n, m in ((20000, 2048), (2048, 1024), (1024, 512), (512, 256)):
    for spec in (1, 0):
        lib.b200_fps_set_speculation(spec)
        us = t(lambda: b200grasp.ops.furthest_point_sampling(x, m))
        print("n=%5d m=%4d speculation=%d: %8.1f us" % (x.shape[1], m, spec, us), flush=True)
    lib.b200_fps_set_speculation(1)
    idx = b200grasp.ops.furthest_point_sampling(x, m)
    x = torch.gather(x, 1, idx.long().unsqueeze(-1).expand(-1, -1, 3)).contiguous()
