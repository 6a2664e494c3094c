#!/usr/bin/env python
"""Throughput of a stream of cfg2 scenes through GraspNetPipeline for several depths (device-resident pool of clouds
larger than L2, CUDA-event timed), and through submit/collect with host clouds."""
import os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import b200grasp
if os.environ.get("B200_ARITH"):
    b200grasp.fused_modules.ARITH = os.environ["B200_ARITH"]
if os.environ.get("B200_STACK_ARITH"):
    b200grasp.fused_modules.STACK_ARITH = os.environ["B200_STACK_ARITH"]
print("arith", b200grasp.fused_modules.ARITH, "stacks", b200grasp.fused_modules.STACK_ARITH)
net = bench.make_net("cuda")
P = 640
host = [bench.scene_cloud(5000 + i) for i in range(P)]
pool = torch.from_numpy(np.stack(host)).cuda()
K = 200
for depth in [int(a) for a in sys.argv[1:]] or (1, 2, 3, 4):
    pipe = b200grasp.graspnet.GraspNetPipeline(net, depth=depth, tiles_per_cta=int(os.environ.get('B200_TPC', '0')) or None)
    for mode in ("device", "host"):
        for rep in range(2):
            torch.cuda.synchronize()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            w0 = time.perf_counter()
            if mode == "device":
                for i in range(K):
                    pipe.submit(device_cloud=pool[i % P], readback=False)
            else:
                pipe.run(host[i % P] for i in range(K))
            pipe.join()
            t1.record()
            torch.cuda.synchronize()
            w1 = time.perf_counter()
        print("depth %d %-6s  %.4f ms/scene (events)  %.4f ms/scene (wall)  %.0f scenes/s" %
              (depth, mode, t0.elapsed_time(t1) / K, (w1 - w0) * 1e3 / K, K / (t0.elapsed_time(t1) * 1e-3)), flush=True)
    del pipe
