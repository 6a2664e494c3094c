#!/usr/bin/env python
"""Timeline (clock64, CTA 0) of the fused group -> MLP -> max-pool kernels: instrumented build
(python graspnet-baseline_b200/build.py --mlp-profile).  argv: crop | sa1 | sa2   [f16 | tf32]"""
import ctypes, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp
from b200grasp import ops, _abi, fused_modules as fm

prof = ctypes.CDLL(os.path.join(ROOT, "graspnet-baseline_b200", "lib", "libb200mlp_prof.so"))
NAMES = ("b200_group_mlp_maxpool", "b200_group_feat_mlp_maxpool", "b200_group_mlp_maxpool_f16", "b200_group_feat_mlp_maxpool_f16")
for name in NAMES:
    f = getattr(prof, name); f.argtypes = _abi.SIGNATURES[name]; f.restype = ctypes.c_int


class Proxy:
    def __getattr__(self, k):
        return getattr(prof, k) if k in NAMES else getattr(_abi.lib, k)


ops._lib = Proxy()
which = sys.argv[1] if len(sys.argv) > 1 else "crop"
arith = sys.argv[2] if len(sys.argv) > 2 else "f16"
f16 = arith == "f16"
xyz = torch.from_numpy(b200grasp.synth.tabletop_cloud(20000, 5)[None]).cuda()
torch.manual_seed(0)
inds = ops.furthest_point_sampling(xyz, 2048)
new_xyz = ops.gather_xyz(xyz, inds, transposed=False)
with torch.no_grad():
    if which == "crop":
        mod = fm.CloudCrop(64, 3).cuda().eval()
        seeds = new_xyz[:, :1024].contiguous()
        rot = torch.from_numpy(b200grasp.synth.seed_rotations(1024, 9).reshape(1, 1024, 9)).cuda()
        idx = ops.cylinder_query_multi(seeds, xyz, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64)
        if f16:
            first, l16 = mod._fused.get_f16()
            run = lambda: ops.group_mlp_maxpool_f16(xyz, seeds, idx, first, l16, rot=rot)
        else:
            layers = mod._fused.get()
            run = lambda: ops.group_mlp_maxpool(xyz, seeds, idx, layers, rot=rot, passes=3)
    elif which == "sa1":
        mod = fm.PointnetSAModuleVotes(npoint=2048, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], normalize_xyz=True).cuda().eval()
        idx = ops.ball_query(new_xyz, xyz, 0.04, 64)
        if f16:
            first, l16 = mod._fused.get_f16()
            run = lambda: ops.group_mlp_maxpool_f16(xyz, new_xyz, idx, first, l16, radius=0.04)
        else:
            layers = mod._fused.get()
            run = lambda: ops.group_mlp_maxpool(xyz, new_xyz, idx, layers, radius=0.04, passes=3)
    else:
        npoint, radius, ns, cin, nin = {"sa2": (1024, 0.1, 32, 128, 2048), "sa3": (512, 0.2, 16, 256, 1024),
                                        "sa4": (256, 0.3, 16, 256, 512)}[which]
        mod = fm.PointnetSAModuleVotes(npoint=npoint, radius=radius, nsample=ns, mlp=[cin, 128, 128, 256], normalize_xyz=True).cuda().eval()
        xyz2 = new_xyz[:, :npoint].contiguous()
        new_xyz = new_xyz[:, :nin].contiguous()
        idx = ops.ball_query(xyz2, new_xyz, radius, ns)
        feat_t = torch.randn(1, nin, cin, device="cuda")
        if f16:
            l16 = mod._fused.get_f16()[1]
            run = lambda: ops.group_feat_mlp_maxpool_f16(new_xyz, xyz2, feat_t, idx, l16, radius=radius)
        else:
            layers = mod._fused.get()
            run = lambda: ops.group_feat_mlp_maxpool(new_xyz, xyz2, feat_t, idx, layers, radius=radius, passes=3)
    for _ in range(3):
        run()
torch.cuda.synchronize()
if f16 and which not in ("crop", "sa1"):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize()
    kp = (ctypes.c_ulonglong * 8)()
    prof.b200_mlp_f16_read_kernel_profile(kp)
    print("kernel (CTA 0): entry -> set-up done %d cycles, -> exit %d cycles; CUDA events around the launch %.1f us"
          % (kp[1] - kp[0], kp[2] - kp[0], e0.elapsed_time(e1) * 1e3))
    print("prologue (thread 0): gathered %d, row bound known %d, source slabs out %d cycles after entry" % (kp[3] - kp[0], kp[4] - kp[0], kp[5] - kp[0]))
    kernel_entry = kp[0]
NG = 2
xyzp = f16 and which in ("crop", "sa1")
buf = (ctypes.c_ulonglong * ((5 if f16 else 3) * 16 * 8))()
assert (prof.b200_mlp_f16_read_profile if f16 else prof.b200_mlp_read_profile)(buf) == 0
a = np.array(buf, dtype=np.int64).reshape(-1, 16, 8)
t0 = a[0, 0, 0]
if xyzp:
    t0 = a[0, 0, 0]
    print(which, "xyz_pipe_kernel, cycles relative to the issuer's first tile")
    print("A0/A1: tile start, acc0 ready, [slab: start, published] x 2, [pool of t-1: acc1 ready, done]")
    print("P1/P2: start, -, -, barrier passed, combine | source done, -, [pool: acc1 ready, done]")
    print("MMA  : -, -, layer-2 slab 0..3 issued at, layer 2 committed, layer 1 (this tile) committed")
    for t in range(3, 9):
        for r, nm in enumerate(("A0 ", "A1 ", "P1 ", "P2 ", "MMA")):
            n = 8
            print("tile %2d %s %s" % (t, nm, " ".join("%7d" % (a[r, t, e] - t0 if a[r, t, e] else 0) for e in range(n))))
    print("cycles per tile (issuer, tiles 3..11): %.0f" % ((a[4, 11, 6] - a[4, 3, 6]) / 8.0))
elif f16:
    print(which, arith, "cycles relative to start; group rows = that group's k-th tile (tile 2k + g): start, gathered (before turn wait), "
          "source slabs out, acc(last-1) ready, mid epilogues out, acc(last) ready, last epilogue read, combine done; "
          "mma rows (per tile): first/last slab ready of the last two layers")
    print("tile 0 start is %d cycles after kernel entry" % (t0 - kernel_entry))
    for k in range(8):
        for g in range(NG):
            print("tile %2d  g%d %s" % (k, g, " ".join("%7d" % (a[g, k, e] - t0) for e in range(8))))
    for it in range(12):
        print("mma tile %2d %s" % (it, " ".join("%7d" % (a[NG, it, e] - t0) for e in range(4))))
    print("cycles per tile (mma, tiles 2..11): %.0f" % ((a[NG, 11, 3] - a[NG, 2, 3]) / 9.0))
    sb = (ctypes.c_ulonglong * 16)()
    prof.b200_mlp_f16_read_slab_profile(sb)
    sb = np.array(sb, dtype=np.int64).reshape(2, 8)
    for r in sb:
        print("slab: entry 0, TMEM read +%d, math done +%d, (slot wait) +%d, slot free +%d, stores issued +%d, proxy fence +%d, arrived +%d"
              % tuple(r[i] - r[0] for i in range(1, 8)))
else:
    names = ["tile start", "L(last-1) acc ready", "E(mid) done", "gather(next) done", "L(last) acc ready", "E(last) done",
             "barrier", "combine done"]
    print(which, "cycles relative to tile 0 start; rows: half 0, half 1 (events: %s); mma: first/last slab ready of the last two layers" % ", ".join(names))
    for it in range(12):
        print("tile %2d  h0 %s" % (it, " ".join("%7d" % (a[0, it, e] - t0) for e in range(8))))
        print("         h1 %s" % " ".join("%7d" % (a[1, it, e] - t0) for e in range(8)))
        print("         mma %s" % " ".join("%7d" % (a[2, it, e] - t0) for e in range(4)))
    print("cycles per tile (tiles 1..11): %.0f" % ((a[0, 11, 0] - a[0, 1, 0]) / 10.0))
