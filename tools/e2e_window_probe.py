#!/usr/bin/env python
"""20-scene windows through GraspNetPipeline (each from an idle GPU, like bench.py's timed regions): device-resident
clouds without / with the result copy, host clouds through run(); host wall time of the submit loop and the collects."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, b200grasp
net = bench.make_net("cuda")
P, K = 64, 20
host = [bench.scene_cloud(5000 + i) for i in range(P)]
pool = torch.from_numpy(np.stack(host)).cuda()
pipe = b200grasp.graspnet.GraspNetPipeline(net, depth=24)

def window(mode, first):
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    t0.record()
    slots = []
    for i in range(first, first + K):
        if mode == "device":
            slots.append(pipe.submit(device_cloud=pool[i % P], readback=False))
        elif mode == "device+d2h":
            slots.append(pipe.submit(device_cloud=pool[i % P], readback=True))
        else:
            slots.append(pipe.submit(host[i % P]))
    w1 = time.perf_counter()
    if mode != "device":
        res = [pipe.collect(k) for k in slots]
    w2 = time.perf_counter()
    pipe.join()
    t1.record()
    torch.cuda.synchronize()
    w3 = time.perf_counter()
    return t0.elapsed_time(t1), (w1 - w0) * 1e3, (w2 - w1) * 1e3, (w3 - w0) * 1e3

for mode in ("device", "device+d2h", "host"):
    r = np.array([window(mode, 20 * w) for w in range(10)][1:])
    m = np.median(r, axis=0)
    print("%-11s events %.3f ms  | host: submits %.3f ms, collects %.3f ms, wall %.3f ms" % (mode, m[0], m[1], m[2], m[3]), flush=True)
