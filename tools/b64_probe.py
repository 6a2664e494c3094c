"""Reproduces bench.py's batch-64 leg step by step, dumping the Python stack if a step takes longer than 50 s."""
import faulthandler, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, b200grasp
from b200grasp.pipeline import GeometryScene
from b200grasp import graspnet
faulthandler.dump_traceback_later(50, exit=True)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
t0 = time.time()
def mark(m):
    torch.cuda.synchronize(); print("[%.1fs] %s" % (time.time() - t0, m), flush=True)
if "geo" in sys.argv:
    sc = GeometryScene(batch=B, num_point=20000, num_grasp=1024, seed=4321, device="cuda")
    mark("scene built"); sc.run(); mark("run"); sc.run_graph(); mark("graph")
net = bench.make_net("cuda")
eng = graspnet.GraspNetEngine(net, batch=B, num_point=20000, device="cuda")
eng.cloud.copy_(torch.from_numpy(np.stack([bench.scene_cloud(4321 + s) for s in range(B)])))
mark("engine built")
with torch.no_grad():
    eng.forward_streams()
mark("forward_streams once")
eng.run_graph(); mark("graph captured + replayed")
eng.run_graph(); mark("replayed")
