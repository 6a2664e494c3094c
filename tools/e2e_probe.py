#!/usr/bin/env python
"""Where the end-to-end step spends its host time: wall clock of every call of bench.py's e2e_step."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
from b200grasp.pipeline import GeometryScene
dev = torch.device("cuda", 0)
sc = GeometryScene(batch=1, num_point=20000, num_grasp=1024, seed=1234, device=dev)
host = sc.host
pin = {k: torch.from_numpy(host[k]).pin_memory() for k in ("cloud", "rot")}
side = torch.cuda.Stream(device=dev)
sc.run_graph(post=False); torch.cuda.synchronize()
acc = {}
def tick(name, t0):
    t1 = time.perf_counter(); acc[name] = acc.get(name, 0.0) + (t1 - t0); return t1
N = 200
for it in range(N + 20):
    if it == 20: acc.clear(); torch.cuda.synchronize(); T0 = time.perf_counter()
    t = time.perf_counter()
    sc.cloud.copy_(pin["cloud"], non_blocking=True); sc.rot.copy_(pin["rot"], non_blocking=True); t = tick("h2d copy_", t)
    out = sc.run_graph(post=False); t = tick("graph launch", t)
    with torch.cuda.stream(side):
        det = b200grasp.ModelFreeCollisionDetector(host["cloud"][0], voxel_size=0.01, device=dev); t = tick("detector ctor", t)
        gg = b200grasp.GraspGroup(host["grasps"][0]); t = tick("GraspGroup", t)
        mask = det.detect(gg, approach_dist=0.05, collision_thresh=0.01); t = tick("detect", t)
        sub = gg[~mask]; t = tick("gg[~mask]", t)
        kept = sub.nms(device=dev); t = tick("nms", t)
    seeds = out["sa2_inds"].cpu(); t = tick("seeds.cpu", t)
torch.cuda.synchronize()
total = (time.perf_counter() - T0) / N * 1e3
print("e2e wall %.3f ms/step" % total)
for k, v in acc.items(): print("  %-14s %7.1f us" % (k, v / N * 1e6))
