#!/usr/bin/env python
"""Per-stage cycle counts of the FPS step (instrumented build: python graspnet-baseline_b200/build.py --fps-profile).

    python tools/fps_prof.py     # per (shape, C, NT): cycles/step in compute, warp-reduce, publish, wait, select
"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp  # noqa: E402  (synthetic clouds only)

lib = ctypes.CDLL(os.path.join(ROOT, "graspnet-baseline_b200", "lib", "libb200fps_prof.so"))
vp, ci = ctypes.c_void_p, ctypes.c_int
lib.b200_furthest_point_sampling.argtypes = [vp, ci, ci, ci, vp, vp]
lib.b200_fps_set_cluster.argtypes = [ci]
lib.b200_fps_set_threads.argtypes = [ci]

names = ["compute", "warp-reduce", "publish", "wait+load", "select", "tail"]
for n, m in ((512, 256), (2048, 1024), (20000, 2048)):
    x = torch.from_numpy(b200grasp.synth.tabletop_cloud(n, 3)[None]).cuda()
    out = torch.empty(1, m, dtype=torch.int32, device="cuda")
    for nt in (128, 256):
        for c in (1, 2, 4, 8, 16):
            if n // c < 512 and c > 1:
                continue
            lib.b200_fps_set_cluster(c)
            lib.b200_fps_set_threads(nt)
            for _ in range(2):
                rc = lib.b200_furthest_point_sampling(x.data_ptr(), 1, n, m, out.data_ptr(), None)
            torch.cuda.synchronize()
            if rc:
                print("n=%d C=%d NT=%d rc=%d" % (n, c, nt, rc))
                continue
            buf = (ctypes.c_ulonglong * 6)()
            lib.b200_fps_read_profile(buf)
            v = [b / (m - 1) for b in buf]
            print("n=%5d m=%4d C=%2d NT=%3d total %6.0f cyc/step : " % (n, m, c, nt, sum(v)) +
                  "  ".join("%s %5.0f" % (nm, t) for nm, t in zip(names, v)), flush=True)
