#!/usr/bin/env python
"""Run FPS once per shape (for ncu captures): python tools/fps_once.py [C] [NT]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
lib = b200grasp._abi.lib
if len(sys.argv) > 1: lib.b200_fps_set_cluster(int(sys.argv[1]))
if len(sys.argv) > 2: lib.b200_fps_set_threads(int(sys.argv[2]))
for n, m in ((512, 256), (2048, 1024), (20000, 2048)):
    x = torch.from_numpy(b200grasp.synth.tabletop_cloud(n, 3)[None]).cuda()
    for _ in range(2):
        b200grasp.ops.furthest_point_sampling(x, m)
    torch.cuda.synchronize()
print("done")
