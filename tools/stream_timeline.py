#!/usr/bin/env python
"""CUPTI kernel records (torch.profiler) of a STREAM of cfg2 scenes through GraspNetPipeline (16 in flight): per kernel
the mean duration inside the stream (stretched by whatever shares the SMs), launches per scene and the share of
sum(durations) -- next to tools/graph_timeline.py's one-scene-alone durations.   python tools/stream_timeline.py [scenes]"""
import os, sys, collections
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import b200grasp
from torch.profiler import profile, ProfilerActivity

K = int(sys.argv[1]) if len(sys.argv) > 1 else 96
net = bench.make_net("cuda")
P = 64
pool = torch.from_numpy(np.stack([bench.scene_cloud(5000 + i) for i in range(P)])).cuda()
pipe = b200grasp.graspnet.GraspNetPipeline(net, depth=16)
for i in range(48):
    pipe.submit(device_cloud=pool[i % P], readback=False)
pipe.join()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(K):
        pipe.submit(device_cloud=pool[i % P], readback=False)
    pipe.join()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
t0 = min(e.time_range.start for e in evs); t1 = max(e.time_range.end for e in evs)
agg = collections.OrderedDict()
for e in sorted(evs, key=lambda e: e.time_range.start):
    a = agg.setdefault(e.name[:90], [0, 0.0])
    a[0] += 1; a[1] += e.time_range.end - e.time_range.start
tot = sum(a[1] for a in agg.values())
print("# %d scenes in %.1f us = %.1f us per scene; sum of kernel durations %.1f us per scene" % (K, t1 - t0, (t1 - t0) / K, tot / K))
print("# %8s %8s %7s  kernel" % ("per scene", "mean us", "share"))
for name, (n, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("  %8.2f %8.1f %6.1f%%  %s" % (n / K, d / n, 100 * d / tot, name))
