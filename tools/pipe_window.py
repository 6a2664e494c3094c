"""Median of 9 windows of K scenes (each window starts from an idle GPU, like bench.py's timed regions) for several
pipeline depths: how much of the gap between --steps 20 and the steady state is pipeline fill / drain."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, b200grasp
net = bench.make_net("cuda")
P = 640
pool = torch.from_numpy(np.stack([bench.scene_cloud(5000 + i) for i in range(P)])).cuda()
for depth in [int(a) for a in sys.argv[1:]] or (6, 8, 12, 16):
    pipe = b200grasp.graspnet.GraspNetPipeline(net, depth=depth, tiles_per_cta=int(os.environ.get('B200_TPC', '0')) or None)
    for K in (20, 200):
        ms = []
        first = 0
        for w in range(10):
            torch.cuda.synchronize()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for i in range(first, first + K):
                pipe.submit(device_cloud=pool[i % P], readback=False)
            pipe.join()
            t1.record()
            torch.cuda.synchronize()
            ms.append(t0.elapsed_time(t1))
            first += K
        med = float(np.median(ms[1:]))
        print("depth %2d  K=%3d  %.3f ms/window  %.4f ms/scene  %.0f scenes/s" % (depth, K, med, med / K, K / (med * 1e-3)), flush=True)
    del pipe
