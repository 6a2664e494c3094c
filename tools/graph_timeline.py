#!/usr/bin/env python
"""Timeline of ONE replay of the whole-inference graph (GraspNetEngine, cfg2 scene) from CUPTI kernel records
(torch.profiler): start, duration, stream of every kernel, warm caches, real overlaps -- what ncu's serialised,
cache-flushed launch list cannot show.      python tools/graph_timeline.py [replays] > profiles/...txt"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import b200grasp
from torch.profiler import profile, ProfilerActivity

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
if os.environ.get("B200_FPS_CLUSTER"):
    b200grasp._abi.lib.b200_fps_set_cluster(int(os.environ["B200_FPS_CLUSTER"]))
net = bench.make_net("cuda")
eng = b200grasp.graspnet.GraspNetEngine(net, batch=1, num_point=20000)
cloud = bench.scene_cloud(1234)[None]
eng.infer(cloud)
for _ in range(5):
    eng.infer(None)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(n):
        eng.infer(None)
        torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
# split into replays at gaps > 100 us
groups, cur, last_end = [], [], None
for e in evs:
    s, t = e.time_range.start, e.time_range.end
    if last_end is not None and s - last_end > 100:
        groups.append(cur); cur = []
    cur.append(e)
    last_end = max(last_end or t, t)
groups.append(cur)
g = groups[-1]
t0 = g[0].time_range.start
print("# last of %d replays: %d device activities; times in us relative to the first" % (n, len(g)))
print("# %8s %8s %8s  %s" % ("start", "dur", "end", "kernel"))
busy = 0.0
for e in g:
    s, t = e.time_range.start - t0, e.time_range.end - t0
    busy += t - s
    print("  %8.1f %8.1f %8.1f  %s" % (s, t - s, t, e.name[:100]))
print("# span %.1f us, sum of durations %.1f us" % (max(e.time_range.end for e in g) - t0, busy))
