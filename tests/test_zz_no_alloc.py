"""b200_grasp_nms_static: the no-allocation variant of grasp NMS (scratch in a file-scope static __device__ array,
include/b200grasp.h) against the existing caller-workspace entry point b200_grasp_nms, through the C ABI.

It has no counterpart in the reference.  Collected last on purpose
(a failure here must not hide the other tests); passes alone in a fresh process and after the
rest of the suite.  Tolerance: NMS keep sets are index work -> bit-exact, like tests/test_gpu_parity.py.
"""
import ctypes
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TT, CR = 0.03, math.cos(30.0 / 180.0 * math.pi)        # GraspGroup.nms() defaults
MAX_M = 2048                                           # B200_NMS_STATIC_MAX_M


def _grasps(b200, m, seed):
    cloud = b200.synth.tabletop_cloud(6000, seed)
    g = b200.synth.synthetic_grasps(cloud, m, seed + 1)
    # a dense patch so that suppression really happens: half the grasps sit on 40 scene points
    rng = np.random.default_rng(seed + 2)
    g[: m // 2, 13:16] = cloud[rng.integers(0, 40, m // 2)]
    return g


def _run_pair(b200, torch, g_host):
    """-> (keep rows via the workspace entry point, via the static entry point), both through ctypes"""
    lib = b200._abi.lib
    m = g_host.shape[0]
    g = torch.from_numpy(g_host).cuda()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    vp = lambda t: ctypes.c_void_p(t.data_ptr())
    keep_a = torch.full((max(m, 1),), -7, dtype=torch.int32, device="cuda")
    keep_b = torch.full((max(m, 1),), -7, dtype=torch.int32, device="cuda")
    n_a = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    n_b = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    ws = torch.empty(int(lib.b200_grasp_nms_workspace(m)), dtype=torch.uint8, device="cuda")
    rc = lib.b200_grasp_nms(vp(g), m, TT, CR, vp(keep_a), vp(n_a), vp(ws), ws.numel(), st)
    assert rc == 0, lib.b200_error_string(rc)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.b200_grasp_nms_static(vp(g), m, TT, CR, vp(keep_b), vp(n_b), st)
    assert rc == 0, lib.b200_error_string(rc)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before          # nothing was taken from torch's allocator either
    ka, kb = int(n_a.item()), int(n_b.item())
    return keep_a[:ka].cpu().numpy(), keep_b[:kb].cpu().numpy()


def test_static_nms_equals_workspace_nms_two_calls_in_a_row(b200, oracle_mod):
    import torch
    # two calls in a row that share the static array: a large problem, then a smaller one that must not see the
    # first call's leftovers (ranks, mask words) -- and the large one again
    first = True
    for m, seed in ((1024, 11), (700, 23), (MAX_M, 5), (1, 9), (1024, 11)):
        g = _grasps(b200, m, seed)
        want, got = _run_pair(b200, torch, g)
        assert got.shape == want.shape and (got == want).all(), "m=%d" % m
        assert 0 < len(got) <= m
        if first:                                            # and both equal the oracle's restatement of nms
            assert (got == oracle_mod.grasp_nms(g)).all()
            assert len(got) < m                              # the patch really suppressed something
            first = False


def test_static_nms_empty_and_limit(b200):
    import torch
    lib = b200._abi.lib
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.zeros((MAX_M + 1, 17), dtype=torch.float32, device="cuda")
    keep = torch.zeros(MAX_M + 1, dtype=torch.int32, device="cuda")
    n = torch.full((1,), 123, dtype=torch.int32, device="cuda")
    vp = lambda t: ctypes.c_void_p(t.data_ptr())
    # beyond the built-in scratch: refused with its own error code and a message that says why; nothing runs
    rc = lib.b200_grasp_nms_static(vp(g), MAX_M + 1, TT, CR, vp(keep), vp(n), st)
    assert rc == -4
    assert b"static scratch" in lib.b200_error_string(rc)
    torch.cuda.synchronize()
    assert int(n.item()) == 123
    with pytest.raises(RuntimeError):
        b200._abi.check(rc, "grasp_nms_static")
    # m = 0: count written, nothing else
    rc = lib.b200_grasp_nms_static(vp(g), 0, TT, CR, vp(keep), vp(n), st)
    assert rc == 0
    torch.cuda.synchronize()
    assert int(n.item()) == 0
    assert lib.b200_grasp_nms_static(vp(g), -1, TT, CR, vp(keep), vp(n), st) == -1
