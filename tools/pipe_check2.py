#!/usr/bin/env python
"""resident submits without host syncs in between (the bench's timed stream): do the slots hold the pool's clouds?"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import b200grasp
mode = sys.argv[1] if len(sys.argv) > 1 else "memcpy"
net = bench.make_net("cuda")
P = 16
host = [bench.scene_cloud(7000 + i) for i in range(P)]
pool = torch.from_numpy(np.stack(host)).cuda()
pipe = b200grasp.graspnet.GraspNetPipeline(net, depth=4)
torch.cuda.synchronize()
for i in range(12):
    if mode == "memcpy":
        pipe.submit(device_cloud=pool[i % P], readback=False)
    else:
        k = pipe.k % 4
        pipe.submit(device_cloud=pool[i % P:i % P + 1].expand(2, -1, -1)[1], readback=False)
pipe.join()
torch.cuda.synchronize()
for k in range(4):
    i = 8 + k
    e = pipe.engines[k]
    print(mode, "slot", k, "cloud equal", bool(torch.equal(e.cloud[0], pool[i % P])), "abs sum %.2f" % float(e.cloud.abs().sum()),
          "nrows", int(e.end_points["grasp_nrows"][0]))
