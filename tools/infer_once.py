#!/usr/bin/env python
"""A few replays of the whole-inference graph (GraspNetEngine, cfg2 scene) -- the command ncu lists the launches of.
    python tools/infer_once.py [replays]          (prints the marker line before the replays start)"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import b200grasp
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
net = bench.make_net("cuda")
eng = b200grasp.graspnet.GraspNetEngine(net, batch=1, num_point=20000)
cloud = bench.scene_cloud(1234)[None]
res = eng.infer(cloud)
torch.cuda.synchronize()
print("captured; replaying", n)
for _ in range(n):
    res = eng.infer(None)
print("ok", res[0]["nrows"], res[0]["kept"].shape)
