#!/usr/bin/env python
"""fp16x3 kernels (csrc/mlp_f16.cu) vs 3xTF32 (csrc/mlp_fused.cu) vs an fp64 statement of the same modules: error
(max |err| / max |want|) and time (CUDA events, L2 flushed)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200grasp
from b200grasp import ops, fused_modules as fm, graspnet

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False


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


def rel(got, want):
    return float((got.double() - want).abs().max() / want.abs().max())


def ref_grouped(xyz, new_xyz, idx, folded, radius=0.0, rot=None, feats=None):
    """fp64: group -> centre -> /radius -> rotate -> [xyz | feats] -> MLP -> max over nsample.  idx (G,B,M,ns)"""
    outs = []
    for g in range(idx.shape[0]):
        B, M, ns = idx.shape[1:]
        ii = idx[g].long().reshape(B, M * ns)
        pts = torch.gather(xyz.double(), 1, ii[..., None].expand(-1, -1, 3)).reshape(B, M, ns, 3)
        d = (pts.float() - new_xyz[:, :, None, :]).double() if True else None
        d = (pts.float() - new_xyz[:, :, None, :])
        if radius > 0:
            d = d / radius
        if rot is not None:
            R = rot.view(B, M, 3, 3)
            d = torch.stack([torch.addcmul(torch.addcmul(d[..., 0] * R[:, :, None, 0, c], d[..., 1], R[:, :, None, 1, c]),
                                           d[..., 2], R[:, :, None, 2, c]) for c in range(3)], dim=-1)
        x = d.double()
        if feats is not None:
            f = torch.gather(feats.double(), 1, ii[..., None].expand(-1, -1, feats.shape[2])).reshape(B, M, ns, -1)
            x = torch.cat([x, f], dim=-1)
        for w, b in folded:
            x = torch.relu(x @ w.double().t() + b.double())
        outs.append(x.max(dim=2).values.permute(0, 2, 1))
    return torch.stack(outs, dim=-1)


def main():
    which = sys.argv[1:] or ["crop", "sa1", "sa2", "sa3", "dense"]
    B = int(os.environ.get("B200_B", "1"))
    xyz = torch.from_numpy(np.stack([b200grasp.synth.tabletop_cloud(20000, 5 + i) for i in range(B)])).cuda()
    torch.manual_seed(0)
    inds = ops.furthest_point_sampling(xyz, 2048)
    new_xyz = ops.gather_xyz(xyz, inds, transposed=False)
    with torch.no_grad():
        if "crop" in which:
            crop = fm.CloudCrop(64, 3).cuda().eval()
            b200grasp.synth.fill_state_dict(crop, 3)
            seeds = new_xyz[:, :1024].contiguous()
            rot = torch.from_numpy(np.stack([b200grasp.synth.seed_rotations(1024, 9 + i) for i in range(B)]).reshape(B, 1024, 9)).cuda()
            idx = ops.cylinder_query_multi(seeds, xyz, rot, 0.05, -0.02, [0.01, 0.02, 0.03, 0.04], 64)
            first, l16 = crop._fused.get_f16()
            l32 = crop._fused.get()
            want = ref_grouped(xyz, seeds, idx, crop.mlps.folded(), rot=rot)
            a = ops.group_mlp_maxpool_f16(xyz, seeds, idx, first, l16, rot=rot)
            b = ops.group_mlp_maxpool(xyz, seeds, idx, l32, rot=rot, passes=3)
            print("crop  err f16x3 %.3e  tf32x3 %.3e   f16 vs tf32 %.3e" % (rel(a, want), rel(b, want), rel(a, b.double())))
            ta = timeit(lambda: ops.group_mlp_maxpool_f16(xyz, seeds, idx, first, l16, rot=rot))
            tb = timeit(lambda: ops.group_mlp_maxpool(xyz, seeds, idx, l32, rot=rot, passes=3))
            print("crop  time f16x3 %.1f us  tf32x3 %.1f us" % (ta, tb), flush=True)
        if "sa1" in which:
            sa1 = fm.PointnetSAModuleVotes(npoint=2048, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], normalize_xyz=True).cuda().eval()
            b200grasp.synth.fill_state_dict(sa1, 4)
            idx1 = ops.ball_query(new_xyz, xyz, 0.04, 64)
            first, l16 = sa1._fused.get_f16()
            l32 = sa1._fused.get()
            want = ref_grouped(xyz, new_xyz, idx1[None], sa1.mlp_module.folded(), radius=0.04)[..., 0]
            a = ops.group_mlp_maxpool_f16(xyz, new_xyz, idx1, first, l16, radius=0.04)
            b = ops.group_mlp_maxpool(xyz, new_xyz, idx1, l32, radius=0.04, passes=3)
            print("sa1   err f16x3 %.3e  tf32x3 %.3e" % (rel(a, want), rel(b, want)))
            ta = timeit(lambda: ops.group_mlp_maxpool_f16(xyz, new_xyz, idx1, first, l16, radius=0.04))
            tb = timeit(lambda: ops.group_mlp_maxpool(xyz, new_xyz, idx1, l32, radius=0.04, passes=3))
            print("sa1   time f16x3 %.1f us  tf32x3 %.1f us" % (ta, tb), flush=True)
        for name, npoint, radius, ns, cin, nin in (("sa2", 1024, 0.1, 32, 128, 2048), ("sa3", 512, 0.2, 16, 256, 1024), ("sa4", 256, 0.3, 16, 256, 512)):
            if name not in which:
                continue
            sa = fm.PointnetSAModuleVotes(npoint=npoint, radius=radius, nsample=ns, mlp=[cin, 128, 128, 256], normalize_xyz=True).cuda().eval()
            b200grasp.synth.fill_state_dict(sa, 5)
            src = new_xyz[:, :nin].contiguous()
            dst = new_xyz[:, :npoint].contiguous()
            idx = ops.ball_query(dst, src, radius, ns)
            feat_t = torch.relu(torch.randn(B, nin, cin, device="cuda")) * 3
            _, l16 = sa._fused.get_f16()
            l32 = sa._fused.get()
            want = ref_grouped(src, dst, idx[None], sa.mlp_module.folded(), radius=radius, feats=feat_t)[..., 0]
            a = ops.group_feat_mlp_maxpool_f16(src, dst, feat_t, idx, l16, radius=radius)
            b = ops.group_feat_mlp_maxpool(src, dst, feat_t, idx, l32, radius=radius, passes=3)
            print("%s   err f16x3 %.3e  tf32x3 %.3e" % (name, rel(a, want), rel(b, want)))
            ta = timeit(lambda: ops.group_feat_mlp_maxpool_f16(src, dst, feat_t, idx, l16, radius=radius))
            tb = timeit(lambda: ops.group_feat_mlp_maxpool(src, dst, feat_t, idx, l32, radius=radius, passes=3))
            print("%s   time f16x3 %.1f us  tf32x3 %.1f us" % (name, ta, tb), flush=True)
        if "dense" in which:
            for name, dims, n, c2 in (("fp2", [512, 256, 256], 1024, 256), ("approach", [256, 256, 302, 302], 1024, 0),
                                      ("head_op", [256, 128, 128, 36], 4096, 0), ("head_tol", [256, 128, 128, 12], 4097, 0)):
                g = torch.Generator(device="cuda").manual_seed(1)
                ws = [(torch.randn(co, ci, device="cuda", generator=g) * (2.0 / ci) ** 0.5, torch.randn(co, device="cuda", generator=g) * 0.1)
                      for ci, co in zip(dims[:-1], dims[1:])]
                x = torch.relu(torch.randn(B, dims[0], n, device="cuda", generator=g)) * 2
                relu_last = name == "fp2"

                class Owner(torch.nn.Module):
                    pass
                fold = graspnet._Folded(Owner(), lambda: ws, relu_last=relu_last)
                x1, x2 = (x[:, :dims[0] - c2].contiguous(), x[:, dims[0] - c2:].contiguous()) if c2 else (x, None)
                want = x.double()
                for i, (w, b) in enumerate(ws):
                    want = torch.einsum("oc,bcn->bon", w.double(), want) + b.double()[None, :, None]
                    if relu_last or i + 1 < len(ws):
                        want = torch.relu(want)
                fm.ARITH = "f16x3"
                a = graspnet._stack(x1, fold, x2=x2)
                fm.ARITH = "tf32x3"
                b = graspnet._stack(x1, fold, x2=x2)
                print("%-8s err f16x3 %.3e  fp32 %.3e  shape %s" % (name, rel(a, want), rel(b, want), tuple(a.shape)))
                fm.ARITH = "f16x3"
                ta = timeit(lambda: graspnet._stack(x1, fold, x2=x2))
                fm.ARITH = "tf32x3"
                tb = timeit(lambda: graspnet._stack(x1, fold, x2=x2))
                fm.ARITH = "f16x3"
                print("%-8s time f16x3 %.1f us (%d launches)  fp32 pointwise %.1f us" % (name, ta, sum(len(g_) for g_ in fold.get_f16()), tb), flush=True)


main()
