#!/usr/bin/env python
"""bench.py -- scenes/sec of GraspNet-baseline inference (geometry hot path + post-processing) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" = one scene of BASELINE.json configs[1]: GraspNet inference (seeded random weights, batch 1,
20000 points -> 2048/1024/512/256 SA points -> 1024 seeds, 300 views / 12 angles / 4 depths),
pred_decode, voxel down-sample of the cloud, collision detection of the decoded grasps and NMS --
graspnet-baseline_b200/graspnet.py::GraspNetEngine, one CUDA graph per scene.  Rank r of N
processes its own scenes, no collective on the data path (scaling = weak).  The geometry ops alone
(graspnet-baseline_b200/pipeline.py, stand-in features instead of the MLPs: the previous headline)
are reported beside it under "geometry_path", with the per-kernel roofline table.

Prints ONE JSON line on rank 0 (see the round contract): value = scenes/s with inputs
resident in HBM; e2e = scenes/s through the reference-facing API with HOST buffers (H2D of
the cloud and grasps, host-side detector construction, D2H of masks/keep-sets inside the
timed region); roofline = dominant kernel's algorithmic bytes / live CUDA-event duration
against MEASURED_PEAKS.json; cpu_baseline = the oracle port of the same step on host cores.
--impl reference times that CPU port alone (the reference has no other CPU implementation of
the pointnet2 ops: SURVEY.md fact 2; oracle/ is the restatement, pinned to the reference).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

def log(msg):
    """progress marker on stderr (stdout carries the one JSON line)"""
    sys.stderr.write("[bench %.1fs] %s\n" % (time.perf_counter() - _T0, msg))
    sys.stderr.flush()


_T0 = time.perf_counter()
METRIC = "scenes/sec (20k pts, 1024 seeds, collision+NMS)"
UNIT = "scenes/s"
WORKLOAD = ("cfg2: GraspNet inference (seeded random weights, eval), batch 1, 20000 pts -> SA 2048/1024/512/256 -> "
            "1024 seeds, 300 views / 12 angles / 4 depths, pred_decode, collision (voxel 0.01, approach 0.05, "
            "thresh 0.01) of the decoded grasps + NMS")
PIPE_DEPTH = 24           # scenes in flight (tools/pipe_window.py, round 2 final code, 20-scene windows / steady state at
                          # depth 16: 3466 / 3890, 20: 3712 / 3976, 24: 3711 / 3949, 32: 3709 / 3959 scenes/s; GraspNetPipeline's
                          # throughput settings on top of that -- 4 tiles per CTA in the few-tile fused-MLP launches, the
                          # sampling chain on 8-CTA clusters, >= 16 tile rounds in the xyz kernels: 3867 / 4201)
POOL = 640                # different clouds the timed stream rotates through (640 x 240 kB = 154 MB > L2)
WEIGHT_SEED = 12          # synth.fill_state_dict seed: every seed passes the objectness test -> 1024 grasps per scene
SA_CFG = [(2048, 0.04, 64, 0), (1024, 0.1, 32, 128), (512, 0.2, 16, 256), (256, 0.3, 16, 256)]
HMAX = [0.01, 0.02, 0.03, 0.04]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def tensor_peak():
    """dense bf16 TFLOP/s measured on this pool's B200s (driver-written); tf32 runs at half the bf16 rate"""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["bf16_tflops"]), "measured"
        except Exception:
            pass
    return 2250.0, "fallback (nominal)"


def fused_mlp_report(sc, timed, dev):
    """SA1's and CloudCrop's group -> shared MLP -> max-pool at config 2 sizes through the fused tcgen05 kernel
    (csrc/mlp_fused.cu), next to the composed path (our group ops + torch SharedMLP + max_pool2d)."""
    import torch
    import torch.nn.functional as F
    from b200grasp import ops, fused_modules as fm
    torch.manual_seed(0)
    B = sc.B
    sa1 = fm.PointnetSAModuleVotes(npoint=2048, radius=0.04, nsample=64, mlp=[0, 64, 64, 128], normalize_xyz=True).to(dev).eval()
    crop = fm.CloudCrop(64, 3).to(dev).eval()
    out = sc.run()
    torch.cuda.synchronize()
    xyz, new_xyz, seeds = sc.cloud, out["sa1_xyz"], out["sa2_xyz"]
    idx1, idxc = out["sa1_idx"], out["crop_idx"]
    rot = sc.rot.view(B, -1, 9)
    rows = {"sa1": B * 2048 * 64, "cloudcrop": B * 1024 * 4 * 64}
    flops = {"sa1": 2 * rows["sa1"] * (3 * 64 + 64 * 64 + 64 * 128),
             "cloudcrop": 2 * rows["cloudcrop"] * (3 * 64 + 64 * 128 + 128 * 256)}
    mma_flops = {"sa1": 2 * rows["sa1"] * (64 * 64 + 64 * 128), "cloudcrop": 2 * rows["cloudcrop"] * (64 * 128 + 128 * 256)}
    peak, which = tensor_peak()
    rep = {"what": "group -> centre/normalise/rotate -> 3 x (1x1 conv + folded BN + ReLU) -> max over nsample, one kernel: "
                   "layer 1 (K=3) on CUDA cores, layers 2-3 tcgen05.mma kind::tf32 with TMEM accumulators; passes=3 is "
                   "3xTF32 (fp32-grade, 1e-5 parity), passes=1 plain TF32 (the reference convs' cuDNN default)",
           "tensor_peak_tf32_tflops": round(peak / 2, 1), "peak_source": which + " bf16 / 2", "cases": []}
    n = 10
    with torch.no_grad():
        for passes in (3, 1):
            l1, lc = sa1._fused.get(), crop._fused.get()
            for name, fn in (("sa1", lambda: ops.group_mlp_maxpool(xyz, new_xyz, idx1, l1, radius=0.04, passes=passes)),
                             ("cloudcrop", lambda: ops.group_mlp_maxpool(xyz, seeds, idxc, lc, rot=rot, passes=passes))):
                ms = timed(fn, n, 3) / n
                tf = mma_flops[name] * passes / (ms * 1e-3) / 1e12
                rep["cases"].append({"module": name, "passes": passes, "ms": round(ms, 4),
                                     "fp32_equivalent_tflops": round(flops[name] / (ms * 1e-3) / 1e12, 1),
                                     "roofline": {"bound": "tensor", "achieved": round(tf, 1), "peak": round(peak / 2, 1),
                                                  "unit": "TFLOP/s", "frac": round(tf / (peak / 2), 4)}})

        # the same two modules in fp16x3 arithmetic (csrc/mlp_f16.cu: xyz_pipe_kernel), the path the engine runs
        f1, l1h = sa1._fused.get_f16()
        fc, lch = crop._fused.get_f16()
        for name, fn in (("sa1", lambda: ops.group_mlp_maxpool_f16(xyz, new_xyz, idx1, f1, l1h, radius=0.04)),
                         ("cloudcrop", lambda: ops.group_mlp_maxpool_f16(xyz, seeds, idxc, fc, lch, rot=rot))):
            ms = timed(fn, n, 3) / n
            tf = mma_flops[name] * 3 / (ms * 1e-3) / 1e12
            rep["cases"].append({"module": name, "arith": "f16x3", "passes": 3, "ms": round(ms, 4),
                                 "fp32_equivalent_tflops": round(flops[name] / (ms * 1e-3) / 1e12, 1),
                                 "roofline": {"bound": "tensor", "achieved": round(tf, 1), "peak": round(peak, 1),
                                              "unit": "TFLOP/s", "frac": round(tf / peak, 4)}})

        def composed_sa1():
            g, _ = sa1.grouper(xyz, new_xyz, None)
            return F.max_pool2d(sa1.mlp_module(g), kernel_size=[1, 64])

        def composed_crop():
            g = [gr(xyz, seeds, sc.rot) for gr in crop.groupers]
            g = torch.stack(g, dim=3).view(B, -1, 1024 * 4, 64)
            return F.max_pool2d(crop.mlps(g), kernel_size=[1, 64])

        def query_only():
            ops.ball_query(new_xyz, xyz, 0.04, 64)
            for h in (0.01, 0.02, 0.03, 0.04):
                ops.cylinder_query(seeds, xyz, rot, 0.05, -0.02, h, 64)

        tq = timed(query_only, n, 3) / n
        rep["composed_ms"] = {"what": "same modules unfused: b200 group ops + torch SharedMLP (cuDNN, TF32 allowed) + "
                                      "max_pool2d; the neighbourhood queries they repeat are timed alone and subtracted",
                              "sa1_plus_cloudcrop": round((timed(composed_sa1, n, 3) + timed(composed_crop, n, 3)) / n - tq, 4)}
    rep["fused_ms_sa1_plus_cloudcrop_3xtf32"] = round(sum(c["ms"] for c in rep["cases"] if c["passes"] == 3 and "arith" not in c), 4)
    rep["fused_ms_sa1_plus_cloudcrop_f16x3"] = round(sum(c["ms"] for c in rep["cases"] if c.get("arith") == "f16x3"), 4)
    return rep


# ---------------------------------------------------------------- CPU port of one step (oracle)

class OracleOps:
    """The functions of pointnet2._ext (bindings.cpp:13-26) on CPU tensors, answered by the C oracle
    (oracle/geom_oracle.c, OpenMP): stands in for b200grasp.ops under the network's composed path, i.e. the reference's
    own op sequence with its CUDA-only ops replaced by their CPU restatement."""

    def __init__(self, oracle, torch):
        self.o, self.t = oracle, torch

    def _np(self, *ts):
        return [np.ascontiguousarray(t.detach().numpy()) for t in ts]

    def furthest_point_sampling(self, xyz, npoint):
        return self.t.from_numpy(self.o.furthest_point_sampling(*self._np(xyz), npoint))

    def gather_points(self, features, idx):
        return self.t.from_numpy(self.o.gather_points(*self._np(features, idx)))

    def ball_query(self, new_xyz, xyz, radius, nsample):
        return self.t.from_numpy(self.o.ball_query(*self._np(new_xyz, xyz), radius, nsample))

    def cylinder_query(self, new_xyz, xyz, rot, radius, hmin, hmax, nsample):
        return self.t.from_numpy(self.o.cylinder_query(*self._np(new_xyz, xyz, rot), radius, hmin, hmax, nsample))

    def group_points(self, features, idx):
        return self.t.from_numpy(self.o.group_points(*self._np(features, idx)))

    def three_nn(self, unknown, known):
        d2, idx = self.o.three_nn(*self._np(unknown, known))
        return self.t.from_numpy(d2), self.t.from_numpy(idx)

    def three_interpolate(self, features, idx, weight):
        return self.t.from_numpy(self.o.three_interpolate(*self._np(features, idx, weight)))


def make_net(device="cpu", passes=3):
    """GraspNet with the configuration of test.py:52-53 and seeded random parameters / BatchNorm statistics"""
    import b200grasp
    net = b200grasp.graspnet.GraspNet(input_feature_dim=0, num_view=300, num_angle=12, num_depth=4, cylinder_radius=0.05,
                                      hmin=-0.02, hmax_list=HMAX, is_training=False, passes=passes)
    b200grasp.synth.fill_state_dict(net, WEIGHT_SEED)
    return net.to(device).eval()


def scene_cloud(seed):
    from b200grasp import synth
    return synth.tabletop_cloud(20000, seed)


def cpu_inference(net, oracle, torch, cloud):
    """One scene on the host: network forward (composed path: reference op sequence, oracle ops + torch CPU layers),
    pred_decode, voxel down-sample, collision mask, NMS -- the step GraspNetEngine runs on the GPU."""
    with torch.no_grad():
        ep = net({"point_clouds": torch.from_numpy(cloud[None])})
    rows = oracle.post.pred_decode(*[ep[k][0].numpy() for k in
                                     ("objectness_score", "grasp_score_pred", "grasp_angle_cls_pred", "grasp_width_pred",
                                      "grasp_tolerance_pred", "fp2_xyz", "grasp_top_view_xyz")])
    scene = oracle.voxel_down_sample(cloud.astype(np.float64), 0.01)
    mask = oracle.collision_detect(scene, rows, 0.01, 0.05, 0.01)
    free = rows[~mask]
    return rows, mask, free[oracle.grasp_nms(free)]


def cpu_modules():
    """The network's module classes for the CPU port WITHOUT loading libb200grasp.so: the package's module files
    (graspnet.py, fused_modules.py, pointnet2_utils.py, synth.py -- the reference-shaped composed path) are imported
    under a private package name whose `_abi` is a stub with no library behind it, and whose pointnet2 `_ext` is the
    C oracle.  Nothing of the CUDA product is mapped into the process that times the CPU baseline."""
    import importlib
    import types
    name = "b200grasp_cpu"
    if name in sys.modules:
        return sys.modules[name]
    pkg_dir = os.path.join(ROOT, "graspnet-baseline_b200")
    pkg = types.ModuleType(name)
    pkg.__path__ = [pkg_dir]
    pkg.__package__ = name
    sys.modules[name] = pkg

    class _NoLib:
        def __getattr__(self, k):
            raise RuntimeError("the CPU baseline must not call the CUDA library (%s)" % k)

    abi = types.ModuleType(name + "._abi")
    abi.lib, abi.LIB_PATH, abi.MlpLayer = _NoLib(), None, None
    abi.check = lambda code, what: None
    sys.modules[name + "._abi"] = abi
    pkg._abi = abi
    for sub in ("ops", "pointnet2_utils", "synth", "fused_modules", "graspnet"):
        setattr(pkg, sub, importlib.import_module(name + "." + sub))
    return pkg


def make_net_cpu(pkg):
    net = pkg.graspnet.GraspNet(input_feature_dim=0, num_view=300, num_angle=12, num_depth=4, cylinder_radius=0.05,
                                hmin=-0.02, hmax_list=HMAX, is_training=False)
    pkg.synth.fill_state_dict(net, WEIGHT_SEED)
    return net.eval()


def time_cpu(steps, warmup):
    # the C port is OpenMP-parallel; torchrun exports OMP_NUM_THREADS=1, which would make this a 1-thread run
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    import torch
    import oracle
    oracle.cpu.build()
    torch.set_num_threads(os.cpu_count() or 1)
    pkg = cpu_modules()
    pkg.fused_modules.FUSED = False
    pkg.pointnet2_utils._ext = OracleOps(oracle, torch)
    net = make_net_cpu(pkg)
    cloud = pkg.synth.tabletop_cloud(20000, 1234)
    for _ in range(warmup):
        cpu_inference(net, oracle, torch, cloud)
    t0 = time.perf_counter()
    for _ in range(steps):
        out = cpu_inference(net, oracle, torch, cloud)
    dt = time.perf_counter() - t0
    return steps / dt, dt / steps, out


# ---------------------------------------------------------------- clocks sampler

class Clocks(threading.Thread):
    """SM clock + throttle reasons sampled WHILE the timed regions run: NVML in-process (a query costs well under
    a millisecond, so even a 10 ms region is seen several times), nvidia-smi as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index, uuid=None):
        super().__init__(daemon=True)
        self.index, self.stop_flag = index, False
        self.sm, self.mx, self.reasons, self.source = [], [], set(), "nvidia-smi"
        self.nvml = self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)) if uuid else "")
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx.append(float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)))
            self.nvml, self.source = pynvml, "nvml"
        except Exception:
            self.nvml = None

    def sample(self):
        if self.nvml is not None:
            n = self.nvml
            self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
            try:
                mask = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle)
            except Exception:
                mask = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
            self.reasons |= {k for k, bit in self.BITS.items() if mask & bit}
            return
        o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                            "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        r = [c.strip() for c in o.strip().split(",")]
        self.sm.append(float(r[0])); self.mx.append(float(r[1]))
        for nme, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[2:6]):
            if v.lower().startswith("active"):
                self.reasons.add(nme)

    def run(self):
        while not self.stop_flag:
            try:
                self.sample()
            except Exception:
                pass
            time.sleep(0.001 if self.nvml is not None else 0.1)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None,
                "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": self.source}


# per-launch DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) from the ncu --set full captures committed
# under profiles/ (a bench run cannot profile itself; the `traffic_source` key of the roofline names the file)
NCU_TRAFFIC = {"fps_sa1": 257536,                        # profiles/r01_top_kernels_full.txt
               # the CloudCrop launch of xyz_pipe_kernel, captured this round: profiles/r02_mlp_f16_ncu.txt
               "group_mlp_maxpool_cloudcrop": 1584384}

# ---------------------------------------------------------------- algorithmic bytes

def alg_bytes(name, B=1, N=20000, M=1024, nscene=7000):
    """Compulsory HBM traffic of one launch: unique inputs read once + outputs written once."""
    npts = [N] + [c[0] for c in SA_CFG]
    if name.startswith("fps_sa"):
        l = int(name[-1]); return B * (npts[l - 1] * 12 + npts[l] * 4)
    if name.startswith("gather_sa"):
        l = int(name[-1]); return B * (npts[l - 1] * 12 + npts[l] * 4 + npts[l] * 12)
    if name.startswith("ball_sa"):
        l = int(name[-1]); ns = SA_CFG[l - 1][2]
        return B * (npts[l - 1] * 12 + npts[l] * 12 + npts[l] * ns * 4)
    if name.startswith("group_xyz_sa"):
        l = int(name[-1]); ns = SA_CFG[l - 1][2]
        return B * (npts[l - 1] * 12 + npts[l] * ns * 4 + 3 * npts[l] * ns * 4)
    if name.startswith("group_feat_sa"):
        l = int(name[-1]); ns, c = SA_CFG[l - 1][2], SA_CFG[l - 1][3]
        return B * (c * npts[l - 1] * 4 + npts[l] * ns * 4 + c * npts[l] * ns * 4)
    if name == "three_nn_fp1": return B * (512 * 12 + 256 * 12 + 512 * 24)
    if name == "three_nn_fp2": return B * (1024 * 12 + 512 * 12 + 1024 * 24)
    if name == "three_interp_fp1": return B * (256 * 256 * 4 + 512 * 24 + 256 * 512 * 4)
    if name == "three_interp_fp2": return B * (256 * 512 * 4 + 1024 * 24 + 256 * 1024 * 4)
    if name == "cylinder_multi": return B * (N * 12 + 1024 * 12 + 1024 * 36 + 4 * 1024 * 64 * 4)
    if name.startswith("cylinder_"): return B * (N * 12 + 1024 * 12 + 1024 * 36 + 1024 * 64 * 4)
    if name == "group_xyz_crop": return B * (N * 12 + 4 * 1024 * 64 * 4 + 3 * 4 * 1024 * 64 * 4)
    if name == "grid_build": return B * (N * 12 + N * 16)      # cloud in, binned {x,y,z,index} out
    if name == "voxel": return N * 24 + nscene * 24
    if name == "collision": return nscene * 24 + M * 68 + M
    if name == "compact": return M * 68 * 2 + M
    if name == "nms": return M * 68 * 2 + 4
    return 0


# ---------------------------------------------------------------- legacy CUDA ops of the reference (same B200)

def legacy_step(ext, sc, torch):
    """The geometry ops of one scene through the reference's OWN CUDA extension (pointnet2._ext built from
    /root/reference into oracle/_ref by oracle/build_ref.py), same inputs and op order as GeometryScene.run_geometry."""
    xyz = sc.cloud
    lv = []
    for lvl, (npoint, radius, nsample, cin) in enumerate(SA_CFG):
        xyz_t = xyz.transpose(1, 2).contiguous()
        inds = ext.furthest_point_sampling(xyz, npoint)
        new_xyz = ext.gather_points(xyz_t, inds).transpose(1, 2).contiguous()
        idx = ext.ball_query(new_xyz, xyz, radius, nsample)
        ext.group_points(xyz_t, idx)
        if cin:
            ext.group_points(sc.feat["sa%d_in" % (lvl + 1)], idx)
        lv.append(new_xyz)
        xyz = new_xyz
    for name, u, k in (("fp1", lv[2], lv[3]), ("fp2", lv[1], lv[2])):
        d2, i3 = ext.three_nn(u, k)
        ext.three_interpolate(sc.feat[name + "_known"], i3, torch.full_like(d2, 1.0 / 3.0))
    rot9 = sc.rot.view(sc.B, -1, 9)
    cloud_t = sc.cloud.transpose(1, 2).contiguous()
    for h in HMAX:
        ci = ext.cylinder_query(lv[1], sc.cloud, rot9, 0.05, -0.02, h, 64)
        ext.group_points(cloud_t, ci)


def load_legacy():
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "pointnet2")):
        return None
    import torch  # noqa: F401
    if ref_dir not in sys.path:
        sys.path.insert(0, ref_dir)
    try:
        import pointnet2._ext as ext
        return ext
    except Exception:
        return None


def inference_op_times(torch, b200grasp, net, cloud, flush, reps=7):
    """Per-op CUDA-event times of one whole scene run call by call on one stream (network forward, decode, device
    post-processing): every public function of b200grasp.ops is wrapped for the duration of the pass; what is left of
    the total is the torch side (fp32 GEMMs of the FP modules / ApproachNet / heads, concatenations, copies)."""
    import inspect
    from b200grasp import ops, graspnet, postprocess
    rec = []

    def wrap(name, fn):
        def w(*a, **k):
            st = torch.cuda.current_stream()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            r = fn(*a, **k)
            e1.record(st)
            rec.append((name, e0, e1))
            return r
        return w

    orig = {n: f for n, f in vars(ops).items() if inspect.isfunction(f) and not n.startswith("_")
            and f.__module__ == ops.__name__}
    per, totals = {}, []
    try:
        for n, f in orig.items():
            setattr(ops, n, wrap(n, f))
        for _ in range(reps + 2):
            flush.zero_()
            del rec[:]
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            with torch.no_grad():
                ep = net({"point_clouds": cloud})
                rows, nrows = b200grasp.pred_decode_device(ep)
                postprocess.grasp_postprocess_device(cloud[0], rows[0], nrows[0:1])
            t1.record()
            torch.cuda.synchronize()
            seen = {}
            for name, e0, e1 in rec:
                k = seen.get(name, 0)
                seen[name] = k + 1
                per.setdefault("%s#%d" % (name, k), []).append(e0.elapsed_time(e1))
            totals.append(t0.elapsed_time(t1))
    finally:
        for n, f in orig.items():
            setattr(ops, n, f)
    rows = [{"op": k, "ms": round(float(np.median(v[2:])), 5)} for k, v in per.items()]
    total = float(np.median(totals[2:]))
    ours = sum(r["ms"] for r in rows)
    rows.sort(key=lambda r: -r["ms"])
    return {"what": "one scene call by call on one stream, eager, L2 flushed first; ops = this library's C-ABI calls in "
                    "call order (#k = k-th call: furthest_point_sampling/gather_xyz/ball_query #0-3 = SA1-SA4, "
                    "group_mlp_maxpool #0 = SA1, #1 = CloudCrop, group_feat_mlp_maxpool #0-2 = SA2-SA4)",
            "total_ms": round(total, 4), "library_ops_ms": round(ours, 4),
            "torch_layers_and_launch_gaps_ms": round(total - ours, 4), "ops": rows}


# ---------------------------------------------------------------- extra legs (rank 0, N = 1)

def nms_leg(torch, b200grasp, timed_local, dev):
    """GraspGroup.nms at M = 1024 with MANY survivors (the headline scene keeps 23 of 1024: its in-graph NMS sees a
    few dozen rows).  Grasps spread over a whole table-top cloud: few pairs are within 3 cm AND 30 degrees."""
    from b200grasp import ops, synth
    cloud = synth.tabletop_cloud(20000, 77)
    out = []
    for name, pts in (("spread over the scene", cloud), ("on 64 scene points (heavy suppression)", cloud[:64])):
        g = torch.from_numpy(synth.synthetic_grasps(pts, 1024, 5)).to(dev)
        keep, nkeep = ops.grasp_nms(g, 0.03, math.cos(30.0 / 180.0 * math.pi))
        n = 20
        ms = timed_local(lambda: ops.grasp_nms(g, 0.03, math.cos(30.0 / 180.0 * math.pi)), n, 3) / n
        out.append({"grasps": 1024, "placement": name, "kept": int(nkeep.item()), "ms": round(ms, 4),
                    "pairs_per_s": round(1024 * 1023 / 2 / (ms * 1e-3), 0)})
    return {"what": "b200_grasp_nms alone (rank + 64x64 bitmask tiles in fp64 + single-CTA sweep), L2 flushed", "cases": out}


def cfg1_leg(torch, b200grasp, dev):
    """BASELINE.json configs[0]: ModelFreeCollisionDetector (voxel 0.01) + detect (approach 0.05, thresh 0.01) +
    GraspGroup.nms on the reference's fixture scene doc/example_data (513,688 valid points -> 7237 voxels; committed
    as tests/golden/scene_example.npz) with 1024 synthetic grasps, through the drop-in classes with HOST arrays in and
    out (wall clock), next to the same calls on the host cores (C port of the numpy path, oracle/post.py)."""
    import oracle
    from b200grasp import synth
    path = os.path.join(ROOT, "tests", "golden", "scene_example.npz")
    if not os.path.exists(path):
        return None
    pts, g = synth.fixture_scene(path)
    sample = pts[g["sample_idx"]].astype(np.float32)
    grasps = synth.synthetic_grasps(sample, 1024, int(g["grasp_seed"]))

    def gpu_step():
        det = b200grasp.ModelFreeCollisionDetector(pts, voxel_size=0.01, device=dev)
        gg = b200grasp.GraspGroup(grasps)
        mask = det.detect(gg, approach_dist=0.05, collision_thresh=0.01)
        return mask, gg[~mask].nms(device=dev)

    def cpu_step():
        scene = oracle.voxel_down_sample(pts, 0.01)
        mask = oracle.collision_detect(scene, grasps, 0.01, 0.05, 0.01)
        return mask, grasps[~mask][oracle.grasp_nms(grasps[~mask])]

    def wall(fn, n):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            out = fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n * 1e3, out

    g_ms, (gm, gk) = wall(gpu_step, 10)
    c_ms, (cm, ck) = wall(cpu_step, 3)
    same = bool(np.array_equal(gm, cm) and np.array_equal(gk.grasp_group_array, ck) and np.array_equal(gm, g["mask_voxel"]))
    return {"workload": "cfg1: detector construction (513688 points -> voxel 0.01 -> %d points) + detect of 1024 grasps + nms, "
                        "host arrays in / out" % int(g["num_voxels"]),
            "gpu_ms": round(g_ms, 3), "cpu_port_ms": round(c_ms, 2), "cpu_cores": os.cpu_count(), "speedup": round(c_ms / g_ms, 1),
            "colliding": int(gm.sum()), "kept_after_nms": int(len(gk)),
            "results_equal_cpu_port_and_reference_detect": same,
            "h2d_bytes": int(pts.nbytes + grasps.shape[0] * 22 * 8 + grasps.nbytes)}


def cfg4_leg(torch, b200grasp, timed_local, dev, peak):
    """BASELINE.json configs[3]: 200k points, FPS to 4096 seeds, 4096 grasps collision-checked against the FULL
    cloud (no voxelisation), NMS of the 4096.  Parity of these shapes: tests/test_gpu_full_size.py."""
    from b200grasp import ops, synth
    n_pts, n_seed = 200000, 4096
    cloud64 = synth.tabletop_cloud(n_pts, 9, extent=0.6, dtype=np.float64)
    cloud = torch.from_numpy(cloud64.astype(np.float32))[None].to(dev)
    g_host = synth.synthetic_grasps(cloud64, n_seed, 2)
    T = torch.from_numpy(np.ascontiguousarray(g_host[:, 13:16], dtype=np.float64)).to(dev)
    R = torch.from_numpy(np.ascontiguousarray(g_host[:, 4:13], dtype=np.float64)).to(dev)
    det = object.__new__(b200grasp.ModelFreeCollisionDetector)      # the detector's own half-space bounds, no scene
    det.finger_width, det.finger_length = 0.01, 0.06                # (utils/collision_detector.py:29-30)
    thr = torch.from_numpy(np.ascontiguousarray(det._thresholds(b200grasp.GraspGroup(g_host), 0.05))).to(dev)
    scene = torch.from_numpy(cloud64).to(dev)
    g = torch.from_numpy(g_host).to(dev)
    n = 5
    fps_ms = timed_local(lambda: ops.furthest_point_sampling(cloud, n_seed), n, 2) / n
    inds = ops.furthest_point_sampling(cloud, n_seed)
    seeds = ops.gather_xyz(cloud, inds, transposed=False)
    grid_ms = timed_local(lambda: ops.query_grid_build(cloud, 0.026), n, 2) / n
    grid = ops.query_grid_build(cloud, 0.026)
    ball_ms = timed_local(lambda: ops.ball_query(seeds, cloud, 0.04, 64, grid=grid), n, 2) / n
    col_ms = timed_local(lambda: ops.collision_counts(scene, T, R, thr), n, 2) / n
    nms_ms = timed_local(lambda: ops.grasp_nms(g, 0.03, math.cos(30.0 / 180.0 * math.pi)), n, 2) / n
    pairs = float(n_pts) * n_seed
    col_bytes = n_pts * 24 + n_seed * (3 + 9 + 10) * 8 + n_seed * 24
    return {"workload": "cfg4: 200000 points (table-top, 1.2 m), FPS -> 4096, grid ball query r=0.04 ns=64 of the 4096 seeds, "
                        "4096 grasps x full cloud collision (fp64), NMS of 4096",
            "fps_ms": round(fps_ms, 4), "fps_ns_per_sample": round(fps_ms * 1e6 / (n_seed - 1), 1),
            "grid_build_ms": round(grid_ms, 4), "ball_query_ms": round(ball_ms, 4),
            "collision_ms": round(col_ms, 4), "collision_pair_tests_per_s": round(pairs / (col_ms * 1e-3), 0),
            "collision_fp64_flops_per_s": round(pairs * 30 / (col_ms * 1e-3), 0),
            "collision_roofline": {"bound": "hbm", "achieved": round(col_bytes / (col_ms * 1e-3) / 1e9, 2), "peak": peak,
                                   "unit": "GB/s", "frac": round(col_bytes / (col_ms * 1e-3) / 1e9 / peak, 5),
                                   "alg_bytes": col_bytes,
                                   "note": "819 M point-grasp tests of ~30 fp64 operations on 5 MB of input: bound by "
                                           "the fp64 pipe, not by HBM (the frac says how far from bandwidth-bound it is)"},
            "nms_ms": round(nms_ms, 4),
            "scene_ms_sum": round(fps_ms + grid_ms + ball_ms + col_ms + nms_ms, 4)}


def training_leg(torch, dist, b200grasp, dev, world, rank):
    """forward (with label generation) + get_loss + backward + Adam at batch 8 per GPU, 20000 points, 5 objects x 300
    annotated grasp points per scene (the dataset has ~1200: 14 GB of labels per GPU; the label kernels' traffic does
    not depend on it).  N > 1: DistributedDataParallel over NCCL; the gradient all-reduce (4 MB) is timed alone and as
    the difference to a step under no_sync().  N = 1: the same step on the reference's legacy CUDA ops + the composed
    label chain beside it."""
    from b200grasp.train_step import TrainStep
    import gc
    ts = TrainStep(dev, batch=8, num_grasp_points=300)
    ms = ts.time(4, 2)
    ms_nosync = ts.time(4, 1, sync=False) if world > 1 else ms
    ar = ts.allreduce_ms()
    # label generation alone: fused kernels vs the composed index_select chain, on the step's own seeds
    from b200grasp import label_generation as lg
    with torch.no_grad():
        ep = dict(ts.labels)
        inds = torch.stack([torch.randperm(20000, device=dev)[:1024] for _ in range(8)])
        ep["fp2_xyz"] = torch.gather(ts.clouds, 1, inds[..., None].expand(-1, -1, 3)).contiguous()
        ep["grasp_top_view_inds"] = torch.randint(0, 300, (8, 1024), device=dev)
        lab = {}
        for name, fused in (("fused", True), ("composed", False)):
            for it in range(3):
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                e = lg.process_grasp_labels(dict(ep), fused=fused)
                lg.match_grasp_view_and_label(e)
                e1.record()
                torch.cuda.synchronize()
                lab[name] = e0.elapsed_time(e1)
            del e
        e = dict(ep)
        e['_time_label_kernels'] = True
        for _ in range(3):
            e = lg.process_grasp_labels(dict(e), fused=True)
        k_ms, k_bytes = e['_label_kernel_ms'], e['_label_kernel_bytes']
        del e
    hbm = peaks()[0]
    out = {"workload": "cfg5: GraspNet training step (forward with label generation + get_loss + backward + Adam), batch 8 "
                       "per GPU, 20000 points, 5 objects x 300 grasp points x (300,12,4) labels per scene; composed path "
                       "(this library's ops with backward kernels + torch layers, BatchNorm on batch statistics), "
                       "cuDNN TF32 allowed (torch default)",
           "n_gpus": world, "ms_per_step": round(ms, 2), "scenes_per_s": round(world * 8 / (ms * 1e-3), 1),
           "ddp": None if world == 1 else {"backend": "nccl", "gradient_bytes": ts.param_bytes,
                                          "ms_per_step_no_sync": round(ms_nosync, 2),
                                          "allreduce_exposed_ms": round(ms - ms_nosync, 3),
                                          "allreduce_alone_ms": round(ar, 4),
                                          "allreduce_share": round(max(ms - ms_nosync, 0.0) / ms, 4)},
           "label_generation_ms_batch8": {k: round(v, 3) for k, v in lab.items()},
           "label_view_stats_roofline": {"bound": "hbm", "kernel": "label_view_stats + label_view_scores (csrc/labels.cu), 8 scenes",
                                          "alg_bytes": int(k_bytes), "ms": round(k_ms, 4),
                                          "achieved": round(k_bytes / (k_ms * 1e-3) / 1e9, 1), "peak": hbm, "unit": "GB/s",
                                          "frac": round(k_bytes / (k_ms * 1e-3) / 1e9 / hbm, 4),
                                          "traffic": 411202000,
                                          "traffic_source": "profiles/r02_hbm_kernels.txt (ncu, this round): 0.40 GB of DRAM "
                                                            "reads + writes per launch -- seeds that share a nearest grasp "
                                                            "point re-read its rows from L2, so the achieved figure is "
                                                            "L2-served bandwidth, not HBM"},
           "label_generation_note": "process_grasp_labels + match_grasp_view_and_label for 8 scenes: csrc/labels.cu "
                                    "(reads 8 x 236 MB of label / offset rows once) vs the reference's index_select / "
                                    "cat / gather chain on the same knn"}
    del ts
    gc.collect()
    torch.cuda.empty_cache()
    if world == 1:
        ext = load_legacy()
        if ext is not None:
            tl = TrainStep(dev, batch=8, num_grasp_points=300, ext=ext, fused_labels=False)
            out["legacy_ops_ms_per_step"] = round(tl.time(2, 1), 2)
            out["speedup_vs_legacy_ops"] = round(out["legacy_ops_ms_per_step"] / ms, 2)
            del tl
            gc.collect()
            torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------- our arm

def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import b200grasp
    from b200grasp.pipeline import GeometryScene

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    sc = GeometryScene(batch=1, num_point=20000, num_grasp=1024, seed=1234 + 1000 * rank, device=dev)
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)      # > 126 MB L2
    stream = torch.cuda.current_stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            flush.zero_(); fn()
        barrier()
        evs = []
        for _ in range(steps):
            flush.zero_()                                   # L2 flush between timed iterations (untimed)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); fn(); e1.record(stream)
            evs.append((e0, e1))
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        # whole job: MAX of the elapsed device time over ranks (each rank ran `steps` scenes of its own shard)
        return b200grasp.sharding.whole_job_throughput(steps, ms, device=dev)[1]

    def timed_local(fn, steps, warmup):
        """this rank only (no barrier, no reduction): per-kernel legs that rank 0 runs alone"""
        for _ in range(warmup):
            flush.zero_(); fn()
        torch.cuda.synchronize()
        evs = []
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); fn(); e1.record(stream)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs)

    from b200grasp import graspnet, synth
    net = make_net(dev)
    eng = graspnet.GraspNetEngine(net, batch=1, num_point=20000, device=dev)
    cloud_host = scene_cloud(1234 + 1000 * rank)
    eng.views(eng.host_in, "in")["cloud"].copy_(torch.from_numpy(cloud_host)[None])
    eng.io_in.copy_(eng.host_in)
    eng.run_graph()                                         # capture (untimed, once)
    sc.run_graph()
    torch.cuda.synchronize()
    try:
        uuid = torch.cuda.get_device_properties(dev).uuid
    except Exception:
        uuid = None
    clocks = Clocks(local_rank, uuid)
    # ---- the headline: a stream of scenes through GraspNetPipeline (PIPE_DEPTH scenes in flight, each its own
    # batch-1 graph on its own stream).  Inputs rotate through a pool of POOL different clouds = 154 MB > 126 MB L2,
    # resident in HBM; the timed region is ONE pair of events around exactly `steps` scenes.
    pipe = graspnet.GraspNetPipeline(net, depth=PIPE_DEPTH, num_point=20000, device=dev)
    pool_host = [scene_cloud(100000 * (rank + 1) + i) for i in range(POOL)]
    pool_host[0] = cloud_host
    pool = torch.from_numpy(np.stack(pool_host)).to(dev)

    WINDOWS = 9

    def timed_stream(run, steps, warmup):
        """WINDOWS back-to-back timed regions of EXACTLY `steps` scenes each (one CUDA-event pair per region, barrier +
        synchronize on both sides, max over ranks); returns (median, min, max) milliseconds per region.  At the
        driver's --steps 20 one region is ~6 ms: a single one is a noisy sample (round 1: 9-13 % spread)."""
        run(warmup, 0)
        ms, first = [], warmup
        for _ in range(WINDOWS):
            barrier()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record(stream)
            run(steps, first)
            pipe.join()
            t1.record(stream)
            barrier()
            ms.append(b200grasp.sharding.whole_job_throughput(steps, t0.elapsed_time(t1), device=dev)[1])
            first += steps
        return float(np.median(ms)), min(ms), max(ms)

    def resident_stream(count, first):
        for i in range(first, first + count):
            pipe.submit(device_cloud=pool[i % POOL], readback=False)

    def host_stream(count, first):                          # e2e: host cloud in, host results out, every scene
        return pipe.run(pool_host[i % POOL] for i in range(first, first + count))

    log("headline stream")
    clocks.start()                                          # samples only while the timed region below runs
    total_ms, total_min, total_max = timed_stream(resident_stream, args.steps, max(args.warmup, PIPE_DEPTH))
    clk = clocks.summary()
    launches = pipe.launches                                # our kernels per scene (library counter at capture)
    # the resident stream really processed the pool's clouds: a slot fed from HBM returns what the same cloud fed
    # from the host returns
    chk = [pipe.collect(pipe.submit(device_cloud=pool[i], readback=True)) for i in (1, 2, 3)]
    ref = host_stream(3, 1)
    for a, b in zip(chk, ref):
        if not (a["nrows"] == b["nrows"] and a["nrows"] > 0 and np.array_equal(a["mask"], b["mask"])
                and np.array_equal(a["kept"], b["kept"])):
            raise SystemExit("bench: device-fed and host-fed scenes of the stream disagree")
    log("e2e stream")
    e2e_ms, e2e_min, e2e_max = timed_stream(host_stream, args.steps, max(args.warmup, PIPE_DEPTH))
    log("latency")
    lat_ms = timed(eng.run_graph, min(args.steps, 50), args.warmup) / min(args.steps, 50)   # one scene alone, L2 flushed
    lat_e2e_ms = timed(lambda: eng.infer(None), min(args.steps, 50), args.warmup) / min(args.steps, 50)

    def call_sequence():
        """the same scene as the reference's scripts run it (test.py:86-96): net(), pred_decode, detector, nms called
        one after the other through the drop-ins, one stream, eager"""
        with torch.no_grad():
            ep = net({"point_clouds": eng.cloud})
            rows = graspnet.pred_decode(ep)[0]
        gg = b200grasp.GraspGroup(rows.cpu().numpy())
        det = b200grasp.ModelFreeCollisionDetector(cloud_host, voxel_size=0.01, device=dev)
        mask = det.detect(gg, approach_dist=0.05, collision_thresh=0.01)
        return gg, mask, gg[~mask].nms(device=dev)

    log("call sequence")
    seq_steps = min(args.steps, 20)
    seq_ms = timed(call_sequence, seq_steps, args.warmup) / seq_steps
    h2d, d2h = eng.host_in.numel(), eng.host_out.numel()
    got = host_stream(1, 0)[0]                              # pool_host[0] is the engine's cloud
    gg, mask, kept = call_sequence()
    if not (got["nrows"] == len(gg) and np.array_equal(got["mask"], mask)
            and np.array_equal(got["kept"], kept.grasp_group_array)):
        raise SystemExit("bench: the engine graph and the drop-in call sequence disagree")
    infer_counts = {"decoded_grasps": int(got["nrows"]), "colliding": int(got["mask"].sum()),
                    "kept_after_nms": int(got["kept"].shape[0])}

    log("geometry path")
    # ================= geometry ops alone (pipeline.GeometryScene: stand-in features instead of the MLPs)
    gsteps = min(args.steps, 20)
    geo_total_ms = timed(sc.run_graph, gsteps, args.warmup)
    geo_launches = sc.launches
    serial_ms = timed(sc.run, gsteps, args.warmup)          # same ops, one stream, eager launches
    sc.glue = 0
    sc.run()
    glue = sc.glue

    # ---- geometry e2e: host buffers in, host results out, every step
    host = sc.host
    pin_in = sc.host_io_in(pinned=True)                     # cloud | seed rotations | grasp rows | row count
    pin_out = torch.zeros(sc.io_out.numel(), dtype=torch.uint8).pin_memory()
    res = sc.io_views(pin_out, "out")                       # seed indices | collision mask | kept rows | kept count
    geo_h2d, geo_d2h = pin_in.numel(), pin_out.numel()

    def e2e_step():
        # one pinned host buffer -> the graph's static inputs -> ONE graph launch (geometry + voxel down-sample +
        # collision mask + compaction + NMS, all device-resident) -> one pinned host buffer of results
        sc.io_in.copy_(pin_in, non_blocking=True)
        sc.run_graph()
        pin_out.copy_(sc.io_out, non_blocking=True)
        torch.cuda.current_stream().synchronize()           # the caller reads the results now
        return res["kept0"][: int(res["nkept0"][0])], res["mask0"]

    geo_e2e_ms = timed(e2e_step, gsteps, args.warmup)

    # the same scene through the numpy-in / numpy-out drop-ins, called one after the other as test.py:94-96 does
    # (detector rebuilt per scene), on a second stream beside the geometry graph: reported for comparison
    side = torch.cuda.Stream(device=dev)

    pin = {k: torch.from_numpy(host[k]).pin_memory() for k in ("cloud", "rot")}

    def dropin_step():
        sc.cloud.copy_(pin["cloud"], non_blocking=True)
        sc.rot.copy_(pin["rot"], non_blocking=True)
        out = sc.run_graph(post=False)
        with torch.cuda.stream(side):
            det = b200grasp.ModelFreeCollisionDetector(host["cloud"][0], voxel_size=0.01, device=dev)
            gg = b200grasp.GraspGroup(host["grasps"][0])
            mask = det.detect(gg, approach_dist=0.05, collision_thresh=0.01)
            kept = gg[~mask].nms(device=dev)
        out["sa2_inds"].cpu()
        return kept, mask

    dropin_ms = timed(dropin_step, gsteps, args.warmup)
    got_k, got_m = e2e_step()
    ref_k, ref_m = dropin_step()
    if not (np.array_equal(got_m.numpy().astype(bool), ref_m) and np.array_equal(got_k.numpy(), ref_k.grasp_group_array)):
        raise SystemExit("bench: device-resident post-processing and the drop-in call sequence disagree")

    # ---- BASELINE.json configs[4]: the training step, batch 8 per GPU, under DDP when N > 1 (all ranks: collective)
    training = None
    if not args.lean and not args.no_training:
        log("training leg")
        training = training_leg(torch, dist, b200grasp, dev, world, rank)
    if rank != 0:
        return None
    log("legacy leg")
    # ---- the reference's legacy CUDA ops on this same GPU (geometry ops only; rank 0)
    legacy = None
    ext = load_legacy() if world == 1 and not args.lean else None
    if ext is not None:
        lsteps = max(3, min(args.steps, 10))
        leg_ms = timed(lambda: legacy_step(ext, sc, torch), lsteps, 3) / lsteps
        geo_ms = timed(lambda: sc.run_graph(post=False), gsteps, args.warmup) / gsteps
        legacy = {"what": "reference pointnet2._ext (votenet-derived kernels, built unmodified for sm_100) on the same B200: "
                          "FPS, gather, ball query, grouping, three_nn, three_interpolate, 4x cylinder query of one cfg2 scene",
                  "legacy_geometry_ms_per_scene": round(leg_ms, 4), "ours_geometry_ms_per_scene": round(geo_ms, 4),
                  "speedup": round(leg_ms / geo_ms, 2) if geo_ms > 0 else None}
        # the whole network forward as the reference runs it: its own op sequence (composed path), its own CUDA
        # extension for the pointnet2 ops, torch / cuDNN layers (TF32 allowed, torch's default), eager launches
        from b200grasp import fused_modules, pointnet2_utils
        fused, pext = fused_modules.FUSED, pointnet2_utils._ext
        fused_modules.FUSED, pointnet2_utils._ext = False, ext

        def legacy_forward():
            with torch.no_grad():
                return net({"point_clouds": eng.cloud})

        try:
            legf_ms = timed(legacy_forward, lsteps, 3) / lsteps
        finally:
            fused_modules.FUSED, pointnet2_utils._ext = fused, pext

        def our_forward():
            with torch.no_grad():
                return net({"point_clouds": eng.cloud})

        ourf_ms = timed(our_forward, gsteps, args.warmup) / gsteps
        legacy.update({"forward_what": "GraspNet forward (no decode / collision / NMS): reference op sequence with its "
                                       "own pointnet2._ext + torch/cuDNN layers, eager  vs  this package's forward "
                                       "(eager, one stream) and the engine graph (whole step incl. post-processing)",
                       "legacy_forward_ms": round(legf_ms, 4), "ours_forward_eager_ms": round(ourf_ms, 4),
                       "ours_whole_step_graph_latency_ms": round(lat_ms, 4),
                       "ours_whole_step_stream_ms_per_scene": round(total_ms / args.steps, 4),
                       "forward_speedup_vs_graph_latency": round(legf_ms / lat_ms, 2)})
        # the reference's END-TO-END scene as its scripts run it (test.py:86-96, demo_pointcloud.py:228-249): forward on
        # its legacy CUDA ops, pred_decode, D2H, ModelFreeCollisionDetector (open3d voxel + numpy detect) and nms on
        # the host.  The host part is timed with the C port of oracle/post.py (the verbatim numpy detect() takes ~1.3 s
        # per scene; the port is what the CPU baseline uses too) -- conservative for the reference.
        import oracle
        oracle.cpu.build()
        rows_h = gg.grasp_group_array.astype(np.float32)
        t0 = time.perf_counter()
        for _ in range(3):
            scene_h = oracle.voxel_down_sample(cloud_host.astype(np.float64), 0.01)
            m_h = oracle.collision_detect(scene_h, rows_h, 0.01, 0.05, 0.01)
            oracle.grasp_nms(rows_h[~m_h])
        post_ms = (time.perf_counter() - t0) / 3 * 1e3
        leg_e2e_ms = legf_ms + post_ms
        legacy.update({"legacy_host_postprocess_ms": round(post_ms, 3), "legacy_e2e_ms_per_scene": round(leg_e2e_ms, 3)})
    log("batch64")
    # ---- config 3 shape: 64 scenes per graph launch on this GPU (throughput mode; extra, not the headline)
    batch64 = None
    if world == 1 and not args.no_batch64 and not args.lean:
        sc64 = GeometryScene(batch=64, num_point=20000, num_grasp=1024, seed=4321, device=dev)
        sc64.run_graph()
        torch.cuda.synchronize()
        b_steps = 5
        b_ms = timed(sc64.run_graph, b_steps, 3) / b_steps
        batch64 = {"workload": "cfg3: 64 scenes of the cfg2 shape in one graph launch on one GPU",
                   "ms_per_64_scenes": round(b_ms, 4), "scenes_per_s": round(64.0 / (b_ms * 1e-3), 1),
                   "gpu_launches_per_step": sc64.launches}
        per64 = {}
        for _ in range(3):
            flush.zero_()
            sc64.timer = []
            sc64.run()
            torch.cuda.synchronize()
            for name, e0, e1 in sc64.timer:
                per64.setdefault(name, []).append(e0.elapsed_time(e1))
            sc64.timer = None
        rows = []
        for name, v in per64.items():
            if name in ("voxel", "collision", "compact", "nms"):
                continue                                    # per-scene launches, same as batch 1
            ms = float(np.median(v))
            ab = alg_bytes(name, B=64)
            rows.append({"kernel": name, "ms": round(ms, 5), "alg_bytes": ab, "gbs": round(ab / (ms * 1e-3) / 1e9, 1)})
        rows.sort(key=lambda k: -k["ms"])
        batch64["kernels"] = rows[:12]
        batch64["peak_gbs"] = peaks()[0]
        del sc64
        eng64 = graspnet.GraspNetEngine(net, batch=64, num_point=20000, device=dev)
        eng64.cloud.copy_(torch.from_numpy(np.stack([scene_cloud(4321 + s) for s in range(64)])))
        eng64.run_graph()
        torch.cuda.synchronize()
        i_ms = timed(eng64.run_graph, b_steps, 3) / b_steps
        batch64["inference"] = {"workload": "cfg3: 64 scenes of whole GraspNet inference + post-processing in one graph "
                                            "launch on one GPU", "ms_per_64_scenes": round(i_ms, 4),
                                "scenes_per_s": round(64.0 / (i_ms * 1e-3), 1), "gpu_launches_per_step": eng64.launches}
        del eng64
    log("fused mlp report + extra legs")
    # ---- SURVEY.md 8(f) f1 (extra, not part of the step): group -> shared MLP -> max-pool fused on the tensor cores
    fused_mlp = fused_mlp_report(sc, timed, dev) if world == 1 and not args.lean else None
    extra = {}
    if world == 1 and not args.lean:
        log("nms leg")
        extra["nms_1024"] = nms_leg(torch, b200grasp, timed_local, dev)
        log("cfg1 leg")
        extra["cfg1"] = cfg1_leg(torch, b200grasp, dev)
        log("cfg4 leg")
        extra["cfg4"] = cfg4_leg(torch, b200grasp, timed_local, dev, peaks()[0])
        log("inference op times")
    infer_ops = inference_op_times(torch, b200grasp, net, eng.cloud, flush) if not args.lean else None
    log("per-kernel pass")
    # ---- per-kernel pass (rank 0): CUDA events around every launch, L2 flushed before each pass
    per = {}
    reps = 9
    for _ in range(reps):
        flush.zero_()
        sc.timer = []
        sc.run()
        torch.cuda.synchronize()
        for name, e0, e1 in sc.timer:
            per.setdefault(name, []).append(e0.elapsed_time(e1))
        sc.timer = None
    peak, which = peaks()
    nscene = int(b200grasp.ops.voxel_down_sample(sc.cloud[0].double(), 0.01)[1].item())   # voxelised scene size
    kern = []
    for name, v in per.items():
        ms = float(np.median(v))
        ab = alg_bytes(name, nscene=nscene)
        kern.append({"kernel": name, "ms": round(ms, 5), "alg_bytes": ab,
                     "gbs": round(ab / (ms * 1e-3) / 1e9, 3) if ms > 0 else None})
    kern.sort(key=lambda k: -k["ms"])
    top = kern[0]
    fps_roof = {"bound": "hbm", "kernel": top["kernel"], "achieved": top["gbs"], "peak": peak, "unit": "GB/s",
                "frac": round(top["gbs"] / peak, 6),
                # dram__bytes_read.sum + dram__bytes_write.sum of the fps_sa1 launch in the committed ncu --set full
                # capture: the cloud is read once, nothing else touches DRAM
                "traffic": NCU_TRAFFIC.get(top["kernel"]), "peak_source": which,
                "note": "largest kernel of a scene's latency chain; a serial sample chain on one cluster (latency-bound, "
                        "state in registers), see DESIGN.md 4.1"}
    fps_steps = sum(c[0] - 1 for c in SA_CFG)
    fps_ms = sum(k["ms"] for k in kern if k["kernel"].startswith("fps_"))
    fps_roof["latency_model"] = {"fps_steps_per_scene": fps_steps, "fps_ms_per_scene": round(fps_ms, 4),
                                 "ns_per_step_sa1": round(1e6 * next(k["ms"] for k in kern if k["kernel"] == "fps_sa1") / (SA_CFG[0][0] - 1), 1)}
    bw = max((k for k in kern if k["gbs"]), key=lambda k: k["gbs"])
    # ---- the dominant kernel of the headline (a stream of scenes is bound by SM-time, not by the chain): CloudCrop's
    # fused group -> MLP -> max-pool, 129 us on all 148 SMs per scene (FPS: 365 us on 16 SMs).  Timed live, alone,
    # with CUDA events on the launching stream, L2 flushed between iterations.
    with torch.no_grad():
        ep = eng.run_graph()
        crop = net.grasp_generator.crop
        seeds, rot9 = ep["fp2_xyz"], ep["grasp_top_view_rot"].reshape(1, -1, 9).contiguous()
        cidx = b200grasp.ops.cylinder_query_multi(seeds, eng.cloud, rot9, crop.cylinder_radius, crop.hmin, crop.hmax_list,
                                                 crop.nsample)
        layers = crop._fused.get()
        n_it = 20
        f16 = crop.passes == 3 and b200grasp.fused_modules.ARITH == "f16x3"
        if f16:
            first, l16 = crop._fused.get_f16()
            run_crop = lambda: b200grasp.ops.group_mlp_maxpool_f16(eng.cloud, seeds, cidx, first, l16, rot=rot9)
        else:
            run_crop = lambda: b200grasp.ops.group_mlp_maxpool(eng.cloud, seeds, cidx, layers, rot=rot9, passes=crop.passes)
        crop_ms = timed_local(run_crop, n_it, 3) / n_it
    crop_rows = cidx.shape[0] * cidx.shape[2] * cidx.shape[3]
    # tensor-core flops actually issued: three products per K step (hi.hi + hi.lo + lo.hi)
    mma_flops = 2.0 * crop_rows * (64 * 128 + 128 * 256) * crop.passes
    tpeak, twhich = tensor_peak()
    if not f16:
        tpeak = tpeak / 2                                                   # kind::tf32 runs at half the f16 / bf16 rate
    tf = mma_flops / (crop_ms * 1e-3) / 1e12
    roof = {"bound": "tensor",
            "kernel": "group_mlp_maxpool (CloudCrop: 262144 rows x 3->64->128->256, %s)" % ("fp16x3, xyz_pipe_kernel" if f16 else "3xTF32"),
            "achieved": round(tf, 1), "peak": round(tpeak, 1), "unit": "TFLOP/s", "frac": round(tf / tpeak, 4),
            "traffic": NCU_TRAFFIC["group_mlp_maxpool_cloudcrop"],
            "traffic_source": "profiles/r02_mlp_f16_ncu.txt (ncu --set full of this launch, this round); algorithmic "
                              "bytes = cloud + seeds + rotations + index lists read once = 1.34 MB, the output (4 MB) "
                              "stays in L2",
            "ms": round(crop_ms, 4),
            "peak_source": twhich + (" bf16 cuBLAS burst; tcgen05 kind::f16 measured at 2156 TFLOP/s issue-bound "
                                     "(profiles/r02_mma_rate.txt)" if f16 else " bf16 / 2 (tf32 runs at half the rate)"),
            "algorithmic": "flops = 2 x rows x (64x128 + 128x256) x 3 products of tcgen05.mma per launch (the K=3 first "
                           "layer runs on the CUDA cores and is not counted); fp32-equivalent useful rate = a third of "
                           "achieved",
            "note": "dominant kernel of the headline stream by SM-time; the largest kernel of a single scene's latency "
                    "is fps_sa1 (latency_kernel)",
            "latency_kernel": fps_roof,
            "best_bandwidth_kernel": {"kernel": bw["kernel"], "achieved": bw["gbs"], "frac": round(bw["gbs"] / peak, 4),
                                      "peak": peak, "unit": "GB/s",
                                      "note": "single launch timed with L2 flushed; B=1 sizes are launch/latency-bound, "
                                              "see DESIGN.md"}}
    value = world * args.steps / (total_ms * 1e-3)
    e2e = world * args.steps / (e2e_ms * 1e-3)
    line = {"metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True,
            "timing": {"windows": WINDOWS, "what": "median of %d back-to-back timed regions of exactly `steps` scenes each; "
                                                    "value = steps / median" % WINDOWS,
                       "window_ms": {"median": round(total_ms, 4), "min": round(total_min, 4), "max": round(total_max, 4)},
                       "scenes_per_s": {"median": round(value, 1), "min": round(world * args.steps / (total_max * 1e-3), 1),
                                        "max": round(world * args.steps / (total_min * 1e-3), 1)}},
            "latency": {"single_scene_ms": round(lat_ms, 4), "single_scene_e2e_ms": round(lat_e2e_ms, 4),
                        "what": "one scene alone through one GraspNetEngine graph, L2 flushed before every iteration; "
                                "e2e = host cloud in (pinned), host results out"},
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (fp64 collision/NMS, int32 indices)",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "scenes_per_step_per_gpu": 1, "weights": "synth.fill_state_dict(seed %d), eval-mode BatchNorm" % WEIGHT_SEED,
                       "grouped_mlp_arithmetic": "fp16x3 on tcgen05 kind::f16 (hi|lo fp16 operands, exact power-of-two scaling, "
                                                 "fp32 accumulate: fp32-grade) for the grouped MLPs and the 1x1-conv stacks "
                                                 "(csrc/mlp_f16.cu)",
                       "execution": "GraspNetPipeline: %d scenes in flight, each a batch-1 CUDA graph (GraspNetEngine, "
                                    "forked over streams by data dependence) on its own stream; throughput settings of the "
                                    "slots: 4 tiles per CTA in the few-tile fused-MLP launches, FPS on 8-CTA clusters (the "
                                    "latency leg uses one engine with the latency settings)" % PIPE_DEPTH,
                       "l2": "inputs rotate through a pool of %d different clouds (%d MB > 126 MB L2) resident in HBM; "
                             "the latency / per-kernel legs flush L2 (512 MiB write) between iterations" %
                             (POOL, POOL * 240000 // 1000000),
                       "single_scene_latency_ms": round(lat_ms, 4),
                       "single_scene_latency_e2e_ms": round(lat_e2e_ms, 4),
                       "scene": infer_counts},
            "clocks": clk, "gpu_launches": launches * args.steps,
            "e2e": {"value": round(e2e, 2), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": round(e2e_ms / args.steps, 4),
                    "window_ms": {"median": round(e2e_ms, 4), "min": round(e2e_min, 4), "max": round(e2e_max, 4)},
                    "api": "GraspNetPipeline.run(clouds): per scene the host cloud is copied into a pinned buffer and from "
                           "there into the slot's static input, one graph launch (network + pred_decode + voxel + "
                           "collision + NMS on the device), one copy of the results (decoded count | mask | kept rows | "
                           "kept count) to pinned host memory, numpy results returned in order; %d scenes in flight" %
                           PIPE_DEPTH,
                    "call_sequence_ms_per_step": round(seq_ms, 4),
                    "call_sequence": "net(end_points), pred_decode, GraspGroup, ModelFreeCollisionDetector(...).detect, "
                                     ".nms() through the drop-ins, one after the other as test.py:86-96 (eager, "
                                     "numpy round trips); results equal the engine's (checked in this run)"},
            "vs_legacy_cuda": None if not legacy or "legacy_e2e_ms_per_scene" not in legacy else {
                "what": "north_star's target (>= 10x the reference legacy-CUDA end-to-end scenes/s on one B200): the "
                        "reference's scene = GraspNet forward on its own pointnet2._ext (built unmodified for sm_100, "
                        "torch / cuDNN layers, TF32 allowed) + host post-processing (C port of detect() + nms: "
                        "conservative, the verbatim numpy is slower), against this package's e2e stream and latency",
                "legacy_e2e_scenes_per_s": round(1e3 / legacy["legacy_e2e_ms_per_scene"], 2),
                "ours_e2e_scenes_per_s": round(e2e, 2),
                "e2e_ratio": round(e2e * legacy["legacy_e2e_ms_per_scene"] / 1e3, 1),
                "latency_ratio": round(legacy["legacy_e2e_ms_per_scene"] / lat_e2e_ms, 1),
                "forward_only_ratio_vs_stream": round(legacy["legacy_forward_ms"] / (total_ms / args.steps), 1)},
            "cfg1": extra.get("cfg1"), "nms_1024": extra.get("nms_1024"), "cfg4": extra.get("cfg4"), "training": training,
            "roofline": roof, "legacy_cuda": legacy, "batch64": batch64, "fused_mlp": fused_mlp,
            "inference_ops": infer_ops,
            "geometry_path": {"what": "the geometry ops + post-processing alone (pipeline.GeometryScene: stand-in "
                                      "features instead of the MLPs, 1024 synthetic grasps), one graph per scene",
                              "scenes_per_s": round(world * gsteps / (geo_total_ms * 1e-3), 2),
                              "ms_per_step": round(geo_total_ms / gsteps, 4),
                              "gpu_launches_per_step": geo_launches, "glue_torch_launches_per_step": glue,
                              "single_stream_eager_ms_per_step": round(serial_ms / gsteps, 4),
                              "e2e_scenes_per_s": round(world * gsteps / (geo_e2e_ms * 1e-3), 2),
                              "e2e_h2d_bytes": int(geo_h2d), "e2e_d2h_bytes": int(geo_d2h),
                              "dropin_calls_ms_per_step": round(dropin_ms / gsteps, 4)},
            "kernels": kern}
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-batch64", action="store_true")
    ap.add_argument("--no-training", action="store_true")
    ap.add_argument("--lean", action="store_true", help="headline, e2e and latency legs only (the command ncu lists "
                                                        "the launches of): no legacy / batch-64 / fused-MLP / per-op legs")
    args = ap.parse_args()
    # stdout carries ONE JSON line: anything a library prints there (e.g. NCCL's version banner) goes to stderr instead
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(out_fd, (json.dumps(obj) + "\n").encode())

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        if rank != 0:
            return 0
        # the driver's --steps / --warmup are honoured (a scene takes ~0.3 s on 16 cores: the default 400 + 16 run in
        # about two minutes); only an absurd request is bounded
        args.steps = min(args.steps, 600)
        sps, sec, _ = time_cpu(args.steps, args.warmup)
        cores = os.cpu_count()
        cb = {"value": round(sps, 4), "unit": UNIT, "cores": cores, "kind": "port",
              "sample": "%d whole scenes of the cfg2 workload: the network's reference op sequence with the pointnet2 ops "
                        "answered by oracle/geom_oracle.c (OpenMP) and the layers by torch CPU (all threads), then "
                        "pred_decode / voxel / collision / NMS restated in oracle/post.py; a CPU PORT (the reference "
                        "has no CPU implementation of its pointnet2 ops), libb200grasp.so is not loaded" % args.steps}
        emit({"metric": METRIC, "value": round(sps, 4), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": round(sec * 1e3, 3), "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None, "dtype": "f32 (fp64 collision/NMS, int32 indices)", "data": "synthetic",
                          "config": {"workload": WORKLOAD}, "impl": "reference", "cpu_baseline": cb,
                          "e2e": {"value": round(sps, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        return 0

    line = run_ours(args, rank, world, local_rank)
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline and not args.lean:
            n = 8
            sps, sec, cpu_out = time_cpu(n, 1)
            line["cpu_baseline"] = {"value": round(sps, 4), "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                    "sample": "%d whole scenes of the same workload (network forward: oracle C ops with "
                                              "OpenMP + torch CPU layers; decode / voxel / collision / NMS: oracle)" % n,
                                    "decoded_grasps": int(cpu_out[0].shape[0]), "kept_after_nms": int(cpu_out[2].shape[0])}
        emit(line)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
