"""Times the fused group -> MLP -> max-pool kernel at GraspNet's sizes against the composed path
(group ops + torch SharedMLP + max_pool2d).  CUDA events, L2 flushed between iterations."""
import sys, os
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200grasp
from b200grasp import ops, fused_modules as fm

def timeit(fn, iters=20, warm=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts))

def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    xyz = torch.from_numpy(np.stack([b200grasp.synth.tabletop_cloud(20000, 5 + i) for i in range(B)])).cuda()
    torch.manual_seed(0)
    sa1 = fm.PointnetSAModuleVotes(npoint=2048, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], normalize_xyz=True).cuda().eval()
    crop = fm.CloudCrop(64, 3).cuda().eval()
    inds = ops.furthest_point_sampling(xyz, 2048)
    new_xyz = ops.gather_xyz(xyz, inds, transposed=False)
    seeds = new_xyz[:, :1024].contiguous()
    rot = torch.from_numpy(np.stack([b200grasp.synth.seed_rotations(1024, 9 + i) for i in range(B)]).reshape(B, 1024, 9)).cuda()
    idx1 = ops.ball_query(new_xyz, xyz, 0.04, 64)
    idxc = ops.cylinder_query_multi(seeds, xyz, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64)
    flops = {"sa1": 2 * B * 2048 * 64 * (3 * 64 + 64 * 64 + 64 * 128), "crop": 2 * B * 1024 * 4 * 64 * (3 * 64 + 64 * 128 + 128 * 256)}
    with torch.no_grad():
        for passes in (3, 1):
            sa1.passes = crop.passes = passes
            l1, lc = sa1._fused.get(), crop._fused.get()
            t1 = timeit(lambda: ops.group_mlp_maxpool(xyz, new_xyz, idx1, l1, radius=0.04, passes=passes))
            tc = timeit(lambda: ops.group_mlp_maxpool(xyz, seeds, idxc, lc, rot=rot, passes=passes))
            print("fused passes=%d  B=%d: SA1 mlp %8.1f us (%6.1f TFLOP/s)   CloudCrop mlp %8.1f us (%6.1f TFLOP/s)" %
                  (passes, B, t1, flops["sa1"] / t1 * 1e-6, tc, flops["crop"] / tc * 1e-6))
    rot33 = rot.view(B, 1024, 3, 3)
    for tf32 in (True, False):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        with torch.no_grad():
            def composed_sa1():
                g, _ = sa1.grouper(xyz, new_xyz, None)
                return torch.nn.functional.max_pool2d(sa1.mlp_module(g), kernel_size=[1, 64])
            def composed_crop():
                g = [gr(xyz, seeds, rot33) for gr in crop.groupers]
                g = torch.stack(g, dim=3).view(B, -1, 1024 * 4, 64)
                return torch.nn.functional.max_pool2d(crop.mlps(g), kernel_size=[1, 64])
            t1, tc = timeit(composed_sa1), timeit(composed_crop)
        print("composed (b200 group ops incl. query + torch SharedMLP, cudnn tf32=%s): SA1 %8.1f us   CloudCrop %8.1f us" % (tf32, t1, tc))
    with torch.no_grad():
        tq1 = timeit(lambda: ops.ball_query(new_xyz, xyz, 0.04, 64))
        tqc = timeit(lambda: ops.cylinder_query_multi(seeds, xyz, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64))
    print("queries alone: ball SA1 %.1f us, cylinder x4 %.1f us" % (tq1, tqc))

main()
