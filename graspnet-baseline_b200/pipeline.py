"""The geometry hot path of ONE GraspNet inference scene, as a sequence of C-ABI calls.

Order and shapes follow the reference forward pass:
  SA1..SA4  (models/backbone.py:30-64 via pointnet2_modules.py:210-272):
            furthest_point_sample -> gather_operation -> ball_query -> grouping_operation(xyz[,features])
  FP1, FP2  (pointnet2_modules.py:371-416): three_nn -> three_interpolate
  CloudCrop (models/modules.py:127-130): cylinder_query x4 hmax -> grouping_operation(xyz)
  post      (demo_pointcloud.py:220-250): ModelFreeCollisionDetector.detect counts -> GraspGroup.nms
The SharedMLP / conv stacks between these ops are cuDNN work outside this path
(SURVEY.md section 2.1 #6); the feature tensors they would produce are seeded
random tensors of the right shape, created once, resident in HBM.
"""
import math

import numpy as np
import torch

from . import ops, synth

# (npoint, radius, nsample, C_in features grouped at this level)  -- models/backbone.py:30-64
SA_CFG = [(2048, 0.04, 64, 0), (1024, 0.1, 32, 128), (512, 0.2, 16, 256), (256, 0.3, 16, 256)]
CYL = dict(radius=0.05, hmin=-0.02, hmax_list=[0.01, 0.02, 0.03, 0.04], nsample=64)   # models/graspnet.py:39,44
# cell edge of the one uniform grid that serves SA1's ball query (bounding radius 0.04) and CloudCrop's cylinder
# query (0.064); any value gives the same lists, 0.026 was the fastest of a sweep (profiles/r01_query_grid.txt)
GRID_CELL = 0.026
POST = dict(approach_dist=0.05, collision_thresh=0.01, translation_thresh=0.03,      # test.py:95, GraspGroup.nms()
            rotation_thresh=30.0 / 180.0 * math.pi)
POST_STREAMS = 16         # streams the per-scene collision + NMS launches of a batch are spread over


class GeometryScene:
    """Device-resident inputs of `batch` scenes + run() = one pass of the hot path."""

    def __init__(self, batch=1, num_point=20000, num_grasp=1024, seed=1234, device="cuda",
                 fused_cylinder=True, sa_cfg=None, voxel_size=0.01):
        self.B, self.N, self.M = batch, num_point, num_grasp
        self.dev = torch.device(device)
        self.fused_cylinder = fused_cylinder
        self.sa_cfg = sa_cfg or SA_CFG
        self.voxel_size = voxel_size
        self.host = self.make_host_inputs(seed)
        self.launches = 0
        self.glue = 0
        self.timer = None
        self._graphs = {}
        self._hold = None
        self._streams = None
        self._post_streams = None
        self._check_streams = None
        g = torch.Generator(device="cpu").manual_seed(seed)
        # stand-ins for MLP outputs (resident, never regenerated)
        self.feat = {}
        npts = [s[0] for s in self.sa_cfg]
        for lvl, (_, _, _, cin) in enumerate(self.sa_cfg):
            if cin:
                self.feat["sa%d_in" % (lvl + 1)] = torch.randn(batch, cin, npts[lvl - 1], generator=g).to(self.dev)
        self.feat["fp1_known"] = torch.randn(batch, 256, npts[3], generator=g).to(self.dev)
        self.feat["fp2_known"] = torch.randn(batch, 256, npts[2], generator=g).to(self.dev)
        self.upload()

    # ---- inputs a caller would hand over from the host
    def make_host_inputs(self, seed):
        h = {"cloud": np.stack([synth.tabletop_cloud(self.N, seed + s) for s in range(self.B)])}
        nseed = self.sa_cfg[1][0]
        h["rot"] = np.stack([synth.seed_rotations(nseed, seed + s) for s in range(self.B)])
        h["grasps"] = [synth.synthetic_grasps(h["cloud"][s], self.M, seed + s) for s in range(self.B)]
        return h

    def upload(self):
        """Host inputs -> static device buffers.  All of a step's inputs live in ONE device allocation (`io_in`) and
        all of its host-bound results in another (`io_out`), so a caller moves each with a single copy:
        io_in  = cloud (B,N,3) f32 | rot (B,S,3,3) f32 | per scene: grasp rows (M,17) f32, row count i32
        io_out = seed indices (B,S) i32 | per scene: collision mask (M) u8, kept rows (M,17) f32, kept count i32
        (every segment 256-byte aligned; `io_views` maps names to typed views of either buffer or of a host copy)."""
        h, d = self.host, self.dev
        B, N, M, S = self.B, self.N, self.M, self.sa_cfg[1][0]
        self._in_layout = [("cloud", np.float32, (B, N, 3)), ("rot", np.float32, tuple(h["rot"].shape))]
        for s in range(B):
            self._in_layout += [("grasps%d" % s, np.float32, (M, 17)), ("ngrasp%d" % s, np.int32, (1,))]
        self._out_layout = [("seeds", np.int32, (B, S))]
        for s in range(B):
            self._out_layout += [("mask%d" % s, np.uint8, (M,)), ("kept%d" % s, np.float32, (M, 17)),
                                 ("nkept%d" % s, np.int32, (1,))]
        self.io_in = torch.zeros(self._io_bytes(self._in_layout), dtype=torch.uint8, device=d)
        self.io_out = torch.zeros(self._io_bytes(self._out_layout), dtype=torch.uint8, device=d)
        vin = self.io_views(self.io_in, "in")
        self.cloud, self.rot = vin["cloud"], vin["rot"]
        self.grasps = [vin["grasps%d" % s] for s in range(B)]
        self.ngrasp = [vin["ngrasp%d" % s] for s in range(B)]
        self._out_views = self.io_views(self.io_out, "out")
        self.io_in.copy_(self.host_io_in())

    @staticmethod
    def _io_bytes(layout):
        off = 0
        for _, dt, shape in layout:
            off += (int(np.prod(shape)) * np.dtype(dt).itemsize + 255) // 256 * 256
        return off

    def io_views(self, buf, which):
        """name -> typed view of `buf` (a uint8 torch tensor on any device, or a uint8 numpy array)"""
        layout = self._in_layout if which == "in" else self._out_layout
        tdt = {np.float32: torch.float32, np.int32: torch.int32, np.uint8: torch.uint8}
        views, off = {}, 0
        for name, dt, shape in layout:
            nb = int(np.prod(shape)) * np.dtype(dt).itemsize
            seg = buf[off: off + nb]
            views[name] = seg.view(tdt[dt]).view(shape) if isinstance(buf, torch.Tensor) else seg.view(dt).reshape(shape)
            off += (nb + 255) // 256 * 256
        return views

    def host_io_in(self, pinned=False):
        """The step's inputs as one host buffer in the io_in layout (a caller fills such a buffer per scene)."""
        buf = torch.zeros(self.io_in.numel(), dtype=torch.uint8)
        if pinned:
            buf = buf.pin_memory()
        v = self.io_views(buf, "in")
        v["cloud"].copy_(torch.from_numpy(self.host["cloud"]))
        v["rot"].copy_(torch.from_numpy(self.host["rot"]))
        for s, g in enumerate(self.host["grasps"]):
            v["grasps%d" % s].copy_(torch.from_numpy(g))
            v["ngrasp%d" % s].fill_(g.shape[0])
        return buf

    def post_scene(self, s):
        """decode's rows of scene s -> voxel down-sample of its cloud, collision mask, surviving rows, NMS: the
        device-resident form of test.py:94-96 + GraspGroup.nms() (postprocess.py)."""
        voxel = self._op('voxel', ops.voxel_down_sample, self.cloud[s].double(), self.voxel_size)
        mask = self._op('collision', ops.collision_mask_device, voxel[0], voxel[1], self.grasps[s], self.ngrasp[s],
                        self.voxel_size, POST["approach_dist"], POST["collision_thresh"])
        free = self._op('compact', ops.compact_rows, self.grasps[s], mask, self.ngrasp[s])
        kept = self._op('nms', ops.grasp_nms_device, free[0], free[1], POST["translation_thresh"],
                        math.cos(POST["rotation_thresh"]))
        return mask, kept

    # ---- optional per-op CUDA-event timing (bench.py's roofline pass)
    def _op(self, name, fn, *args, **kw):
        if self.timer is None:
            return fn(*args, **kw)
        st = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        r = fn(*args, **kw)
        e1.record(st)
        self.timer.append((name, e0, e1))
        return r

    # ---- one pass, as a CUDA graph over several streams
    def run_graph(self, post=True):
        """One pass as a single graph launch.  The pass is captured once: the FPS chain of the four SA
        levels (which only ever occupies a cluster's worth of SMs) runs on its own stream and every
        other op starts on a side stream as soon as the level it needs exists, so the pass costs
        about the FPS chain instead of the sum of all kernels.  Inputs (self.cloud, self.rot,
        self.grasps, ...) are static device buffers: refresh them with copy_() and replay."""
        if post not in self._graphs:
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):              # lazy one-time setup must not be captured
                self.run_streams(post)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            from . import _abi
            n0 = _abi.lib.b200_launch_count()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                out = self.run_streams(post)
            self._graphs[post] = (g, out, int(_abi.lib.b200_launch_count() - n0), self._hold)
        g, out, nl, _ = self._graphs[post]
        g.replay()
        self.launches = nl
        return out

    def run_streams(self, post=True):
        """The same ops as run(), forked over streams by data dependence (joins back before returning)."""
        main = torch.cuda.current_stream()
        if self._streams is None:
            self._streams = [torch.cuda.Stream(device=self.dev) for _ in range(4)]
        s_q, s_c, s_f, s_p = self._streams

        def after(stream, *deps):
            for d in deps:
                stream.wait_stream(d)

        out = {}
        B = self.B
        # every temporary stays referenced until the streams have joined: a block freed on the
        # stream that allocated it could otherwise be handed out again while a side stream still reads it
        hold = []
        # post-processing: independent of the backbone; per-scene launches (the NMS sweep is a one-CTA
        # serial chain) are dealt round-robin over a few streams so that scenes of a batch overlap
        masks, keeps = [None] * B, [None] * B
        if post:
            if self._post_streams is None:
                self._post_streams = [s_p] + [torch.cuda.Stream(device=self.dev) for _ in range(min(B, POST_STREAMS) - 1)]
            for sp in self._post_streams:
                after(sp, main)
            for s in range(B):
                with torch.cuda.stream(self._post_streams[s % len(self._post_streams)]):
                    masks[s], keeps[s] = self.post_scene(s)
                    ov = self._out_views              # host-bound results into the one output buffer
                    ov["mask%d" % s].copy_(masks[s])
                    ov["kept%d" % s].copy_(keeps[s][0])
                    ov["nkept%d" % s].copy_(keeps[s][1])
            out["collision_mask"], out["kept"] = masks, keeps
            self.glue += 1 + 3 * B
        after(s_c, main)
        with torch.cuda.stream(s_c):
            cloud_t = self.cloud.transpose(1, 2).contiguous()
            # one binning of the full cloud serves SA1's ball query and CloudCrop's cylinder query; it runs
            # while the main stream is busy with SA1's FPS
            grid = ops.query_grid_build(self.cloud, GRID_CELL) if self.cloud.shape[1] >= ops.GRID_QUERY_MIN_N else None
            hold.append(grid)
        xyz, xyz_t = self.cloud, cloud_t
        npts = [c[0] for c in self.sa_cfg]
        nested = all(npts[l + 1] <= npts[l] for l in range(len(npts) - 1)) and len(npts) <= 4
        verdicts = []
        for lvl, (npoint, radius, nsample, cin) in enumerate(self.sa_cfg):
            name = "sa%d" % (lvl + 1)
            # main stream = the serial chain: FPS -> gather -> next level's FPS
            if lvl == 0:
                inds = ops.furthest_point_sampling(xyz, npoint)
            elif nested:
                if lvl == 1:
                    after(main, *self._check_streams)
                inds = ops.furthest_point_sampling_spec(xyz, npoint, verdicts[:lvl])
            else:
                inds = ops.furthest_point_sampling(xyz, npoint)
            new_xyz, new_xyz_t = ops.gather_xyz(xyz, inds)
            hold += [xyz_t, new_xyz_t]
            out[name + "_inds"], out[name + "_xyz"] = inds, new_xyz
            if lvl == 0 and nested:
                # SA2..SA4 sample prefixes of SA1's output (if each level returns the identity): all their
                # prefix checks run now, side by side, instead of one after the other on the serial chain
                if self._check_streams is None:
                    self._check_streams = [torch.cuda.Stream(device=self.dev) for _ in range(len(npts) - 1)]
                for l in range(1, len(npts)):
                    cs = self._check_streams[l - 1]
                    after(cs, main)
                    with torch.cuda.stream(cs):
                        verdicts.append(ops.fps_prefix_check(new_xyz, npts[l - 1], npts[l]))
                hold += verdicts
            after(s_q, main, *([s_c] if lvl == 0 else []))
            with torch.cuda.stream(s_q):              # neighbourhood queries + grouping of this level
                idx = ops.ball_query(new_xyz, xyz, radius, nsample, grid=grid if lvl == 0 else None)
                out[name + "_idx"] = idx
                out[name + "_grouped_xyz"] = ops.group_points(xyz_t, idx)
                if cin:
                    out[name + "_grouped_feat"] = ops.group_points(self.feat[name + "_in"], idx)
            if lvl == 1:                              # seeds exist: CloudCrop's cylinder queries
                after(s_c, main)
                with torch.cuda.stream(s_c):
                    rot9 = self.rot.view(B, -1, 9)
                    cidx = ops.cylinder_query_multi(new_xyz, self.cloud, rot9, CYL["radius"], CYL["hmin"],
                                                    CYL["hmax_list"], CYL["nsample"], grid=grid)
                    out["crop_idx"] = cidx
                    if post:
                        self._out_views["seeds"].copy_(inds)
                    cat = cidx.permute(1, 0, 2, 3).reshape(B, cidx.shape[0] * cidx.shape[2], cidx.shape[3])
                    if B > 1:
                        cat = cat.contiguous()
                    hold += [cat, cloud_t]
                    out["crop_grouped_xyz"] = ops.group_points(cloud_t, cat)
            xyz, xyz_t = new_xyz, new_xyz_t
        after(s_f, main)
        with torch.cuda.stream(s_f):                  # feature propagation
            for name, unknown, known in (("fp1", out["sa3_xyz"], out["sa4_xyz"]), ("fp2", out["sa2_xyz"], out["sa3_xyz"])):
                d2, idx3 = ops.three_nn(unknown, known)
                w = torch.full_like(d2, 1.0 / 3.0)
                hold += [d2, w]
                out[name + "_idx"] = idx3
                out[name + "_interp"] = ops.three_interpolate(self.feat[name + "_known"], idx3, w)
        after(main, s_q, s_c, s_f, *(self._post_streams if post else []))
        self._hold = hold
        return out

    # ---- one pass, one stream
    def run(self):
        out = self.run_geometry()
        L = self.launches
        masks, keeps = [], []
        for s in range(self.B):
            mask, kept = self.post_scene(s)
            masks.append(mask)
            keeps.append(kept)
            L += 7 + 3 + 1 + 4
        out["collision_mask"], out["kept"] = masks, keeps
        self.launches = L
        return out

    def run_geometry(self):
        out = {}
        L = 0
        xyz = self.cloud
        grid = None
        if self.cloud.shape[1] >= ops.GRID_QUERY_MIN_N:
            grid = self._op('grid_build', ops.query_grid_build, self.cloud, GRID_CELL)
            L += 4
        for lvl, (npoint, radius, nsample, cin) in enumerate(self.sa_cfg):
            xyz_t = xyz.transpose(1, 2).contiguous(); self.glue += 1
            inds = self._op('fps_sa%d' % (lvl + 1), ops.furthest_point_sampling, xyz, npoint)
            new_xyz = self._op('gather_sa%d' % (lvl + 1), ops.gather_points, xyz_t, inds).transpose(1, 2).contiguous(); self.glue += 1
            idx = self._op('ball_sa%d' % (lvl + 1), ops.ball_query, new_xyz, xyz, radius, nsample,
                           grid=grid if lvl == 0 else None)
            out["sa%d_grouped_xyz" % (lvl + 1)] = self._op('group_xyz_sa%d' % (lvl + 1), ops.group_points, xyz_t, idx)
            L += 4
            if cin:
                out["sa%d_grouped_feat" % (lvl + 1)] = self._op('group_feat_sa%d' % (lvl + 1), ops.group_points, self.feat["sa%d_in" % (lvl + 1)], idx)
                L += 1
            out["sa%d_inds" % (lvl + 1)], out["sa%d_xyz" % (lvl + 1)], out["sa%d_idx" % (lvl + 1)] = inds, new_xyz, idx
            xyz = new_xyz
        for name, unknown, known in (("fp1", out["sa3_xyz"], out["sa4_xyz"]), ("fp2", out["sa2_xyz"], out["sa3_xyz"])):
            d2, idx3 = self._op('three_nn_' + name, ops.three_nn, unknown, known)
            w = torch.full_like(d2, 1.0 / 3.0)            # weights are caller arithmetic (pointnet2_modules.py:395-397)
            self.glue += 1
            out[name + "_idx"] = idx3
            out[name + "_interp"] = self._op('three_interp_' + name, ops.three_interpolate, self.feat[name + "_known"], idx3, w)
            L += 2
        seed_xyz = out["sa2_xyz"]
        rot9 = self.rot.view(self.B, -1, 9)
        cloud_t = self.cloud.transpose(1, 2).contiguous(); self.glue += 1
        if self.fused_cylinder:
            cidx = self._op('cylinder_multi', ops.cylinder_query_multi, seed_xyz, self.cloud, rot9, CYL["radius"],
                            CYL["hmin"], CYL["hmax_list"], CYL["nsample"], grid=grid)
            L += 1
        else:
            cidx = torch.stack([self._op('cylinder_%g' % h, ops.cylinder_query, seed_xyz, self.cloud, rot9, CYL["radius"],
                                         CYL["hmin"], h, CYL["nsample"], grid=grid) for h in CYL["hmax_list"]])
            L += len(CYL["hmax_list"]); self.glue += 1
        out["crop_idx"] = cidx
        nh = cidx.shape[0]
        # (H,B,M,ns) -> per batch one (B, H*M, ns) index tensor = one grouping launch for all depths
        cat = cidx.permute(1, 0, 2, 3).reshape(self.B, nh * cidx.shape[2], cidx.shape[3])
        if self.B > 1:
            cat = cat.contiguous(); self.glue += 1
        out["crop_grouped_xyz"] = self._op('group_xyz_crop', ops.group_points, cloud_t, cat)
        L += 1
        self.launches = L
        return out
