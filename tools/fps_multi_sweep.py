#!/usr/bin/env python
"""fps_multi.cu: time per sample vs cloud size and cluster size (NT=128), table-top clouds."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
lib = b200grasp._abi.lib

def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3

for b in (1, 8, 64):
    for n in (3500, 4096, 6000, 8192, 12000, 16384, 20000, 30000, 49000, 98000):
        if b > 1 and n not in (4096, 20000): continue
        m = min(2048, n // 4)
        x = torch.from_numpy(np.stack([b200grasp.synth.tabletop_cloud(n, 17 + i) for i in range(b)])).cuda()
        row = []
        for c in (0, 2, 4, 8, 16):
            lib.b200_fps_set_cluster(c)
            lib.b200_fps_set_threads(128 if c else 0)
            try:
                us = t(lambda: b200grasp.ops.furthest_point_sampling(x, m))
                row.append("C=%2d %7.1f" % (c, us))
            except RuntimeError as e:
                row.append("C=%2d    fail" % c)
        lib.b200_fps_set_multi(0); lib.b200_fps_set_cluster(0); lib.b200_fps_set_threads(0)
        us = t(lambda: b200grasp.ops.furthest_point_sampling(x, m))
        lib.b200_fps_set_multi(1)
        print("b=%2d n=%6d m=%4d us: %s | chain(auto) %7.1f" % (b, n, m, "  ".join(row), us), flush=True)
lib.b200_fps_set_cluster(0); lib.b200_fps_set_threads(0)
