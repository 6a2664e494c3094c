#!/usr/bin/env python
"""One eager pass of the whole hot path (cfg2 scene) -- the command ncu captures the top kernels from."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from b200grasp.pipeline import GeometryScene
sc = GeometryScene(batch=1, num_point=20000, num_grasp=1024, seed=1234, device="cuda")
for _ in range(2):
    sc.run()
torch.cuda.synchronize()
print("ok")
