#!/usr/bin/env python
"""Batch-64 graph time vs number of post-processing streams."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
from b200grasp import pipeline
for ps in (8, 16, 32, 64):
    pipeline.POST_STREAMS = ps
    sc = pipeline.GeometryScene(batch=64, num_point=20000, num_grasp=1024, seed=4321, device="cuda")
    for post in (True, False):
        for _ in range(3): sc.run_graph(post=post)
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); sc.run_graph(post=post); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        print("post streams %2d post=%d: %.3f ms per 64 scenes" % (ps, post, float(np.median(ts))), flush=True)
    del sc
