#!/usr/bin/env python
"""device-resident submits vs host submits of the same clouds through GraspNetPipeline: same results?"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import b200grasp
net = bench.make_net("cuda")
host = [bench.scene_cloud(7000 + i) for i in range(6)]
pool = torch.from_numpy(np.stack(host)).cuda()
pipe = b200grasp.graspnet.GraspNetPipeline(net, depth=3)
want = pipe.run(host)
for i in range(6):
    k = pipe.submit(device_cloud=pool[i], readback=True)
    got = pipe.collect(k)
    eng = pipe.engines[k]
    print(i, "cloud equal", bool(torch.equal(eng.cloud[0], pool[i])), "abs sum %.3f" % float(eng.cloud.abs().sum()),
          "nrows", got["nrows"], want[i]["nrows"], "kept equal", got["kept"].shape == want[i]["kept"].shape and bool((got["kept"] == want[i]["kept"]).all()))
