#!/usr/bin/env python
"""fps_multi.cu (several samples per cluster exchange) vs the chain kernel of fps.cu vs the oracle: equality and time."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp, oracle
lib = b200grasp._abi.lib
oracle.cpu.build()

def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3

for kind, b, n, m in (("table", 1, 20000, 2048), ("adv", 2, 20000, 2048), ("table", 1, 12000, 1024), ("adv", 1, 5000, 5000),
                      ("table", 3, 8192, 4096), ("table", 1, 49000, 1024), ("adv", 1, 3100, 3000), ("table", 64, 20000, 2048)):
    f = b200grasp.synth.tabletop_cloud if kind == "table" else b200grasp.synth.adversarial_cloud
    x = np.stack([f(n, 17 + i) for i in range(b)])
    xd = torch.from_numpy(x).cuda()
    want = oracle.furthest_point_sampling(x[:2], m)
    for c, nt in ((0, 0), (4, 128), (8, 128), (16, 128), (2, 256), (4, 256), (8, 256)):
        lib.b200_fps_set_cluster(c)
        lib.b200_fps_set_threads(nt)
        res = {}
        for multi in (1, 0):
            lib.b200_fps_set_multi(multi)
            got = b200grasp.ops.furthest_point_sampling(xd, m)
            torch.cuda.synchronize()
            ok = (got[:2].cpu().numpy() == want).all()
            us = t(lambda: b200grasp.ops.furthest_point_sampling(xd, m))
            res[multi] = (got, ok, us)
        same = bool((res[1][0] == res[0][0]).all())
        print("%-5s b=%2d n=%6d m=%5d C=%2d NT=%3d: multi %8.1f us (%6.1f ns/sample) %s | chain %8.1f us %s | equal=%s" % (
            kind, b, n, m, c, nt, res[1][2], res[1][2] * 1e3 / (m - 1), "ok" if res[1][1] else "MISMATCH",
            res[0][2], "ok" if res[0][1] else "MISMATCH", same), flush=True)
lib.b200_fps_set_cluster(0); lib.b200_fps_set_threads(0); lib.b200_fps_set_multi(1)
