#!/usr/bin/env python
"""Per-stage cycles of fps_multi.cu (instrumented build: python graspnet-baseline_b200/build.py --fps-profile)."""
import ctypes, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
lib = ctypes.CDLL(os.path.join(ROOT, "graspnet-baseline_b200", "lib", "libb200fps_prof.so"))
vp, ci = ctypes.c_void_p, ctypes.c_int
lib.b200_furthest_point_sampling.argtypes = [vp, ci, ci, ci, vp, vp]
lib.b200_fps_set_cluster.argtypes = [ci]
lib.b200_fps_set_threads.argtypes = [ci]
names = ["compute", "warp-reduce", "publish", "wait+load", "chain", "tail"]
for kind, n, m, c, nt in (("table", 20000, 2048, 8, 128), ("table", 20000, 2048, 16, 128), ("adv", 20000, 2048, 8, 128),
                          ("table", 12000, 1024, 8, 128), ("table", 49000, 1024, 16, 128)):
    f = b200grasp.synth.tabletop_cloud if kind == "table" else b200grasp.synth.adversarial_cloud
    x = torch.from_numpy(f(n, 3)[None]).cuda()
    out = torch.empty(1, m, dtype=torch.int32, device="cuda")
    lib.b200_fps_set_cluster(c)
    lib.b200_fps_set_threads(nt)
    for _ in range(2):
        rc = lib.b200_furthest_point_sampling(x.data_ptr(), 1, n, m, out.data_ptr(), None)
    torch.cuda.synchronize()
    buf = (ctypes.c_ulonglong * 7)()
    lib.b200_fps_multi_read_profile(buf)
    rounds = max(1, buf[6])
    print("%s n=%d m=%d C=%d NT=%d rc=%d: rounds %d (%.2f samples/round), total %.0f cycles, %.0f/sample" % (
        kind, n, m, c, nt, rc, rounds, (m - 1) / rounds, sum(buf[:6]), sum(buf[:6]) / (m - 1)))
    print("   per round : " + "  ".join("%s %5.0f" % (nm, b / rounds) for nm, b in zip(names, buf)))
    print("   per sample: " + "  ".join("%s %5.0f" % (nm, b / (m - 1)) for nm, b in zip(names, buf)), flush=True)
