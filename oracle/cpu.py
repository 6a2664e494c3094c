"""ctypes front-end of oracle/geom_oracle.c plus the numpy-side restatements.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Every function cites the
reference file:line it restates (paths relative to /root/reference).
"""
import ctypes
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "geom_oracle.c")
_SO = os.path.join(_HERE, "libgeom_oracle.so")
_lib = None


def build(force=False):
    """gcc the C restatement (-ffp-contract=off: the file states every fma itself)."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC",
                               "-o", _SO, _SRC, "-lm"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.oracle_grasp_nms_opts.restype = ctypes.c_int
        _lib.oracle_grasp_nms_opts.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                               ctypes.c_uint, ctypes.c_void_p]
        _lib.oracle_grasp_nms.restype = ctypes.c_int
        _lib.oracle_grasp_nms.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                          ctypes.c_double, ctypes.c_void_p]
        _lib.oracle_opt_n_threads.restype = ctypes.c_int
    return _lib


def _p(a):
    return ctypes.c_void_p(a.ctypes.data)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _ci(v):
    return ctypes.c_int(int(v))


def _cf(v):
    return ctypes.c_float(float(v))


# --- pointnet2 ops -----------------------------------------------------------

def furthest_point_sampling(xyz, npoint, tree=False):
    """sampling_gpu.cu:74-178 (+ sampling.cpp:70-91). xyz (B,N,3) f32 -> (B,npoint) i32."""
    xyz = _f32(xyz)
    b, n, _ = xyz.shape
    out = np.zeros((b, npoint), np.int32)
    fn = lib().oracle_fps_tree if tree else lib().oracle_fps
    fn(_p(xyz), _ci(b), _ci(n), _ci(npoint), _p(out))
    return out


def gather_points(points, idx):
    """sampling_gpu.cu:13-25. points (B,C,N), idx (B,M) -> (B,C,M)."""
    points, idx = _f32(points), _i32(idx)
    b, c, n = points.shape
    m = idx.shape[1]
    out = np.empty((b, c, m), np.float32)
    lib().oracle_gather_points(_p(points), _p(idx), _ci(b), _ci(c), _ci(n), _ci(m), _p(out))
    return out


def gather_points_grad(grad_out, idx, n):
    """sampling_gpu.cu:39-52. grad_out (B,C,M), idx (B,M) -> (B,C,n)."""
    grad_out, idx = _f32(grad_out), _i32(idx)
    b, c, m = grad_out.shape
    out = np.empty((b, c, n), np.float32)
    lib().oracle_gather_points_grad(_p(grad_out), _p(idx), _ci(b), _ci(c), _ci(n), _ci(m), _p(out))
    return out


def ball_query(new_xyz, xyz, radius, nsample):
    """ball_query_gpu.cu:14-49 (argument order of the pybind entry, ball_query.cpp:13)."""
    new_xyz, xyz = _f32(new_xyz), _f32(xyz)
    b, m, _ = new_xyz.shape
    n = xyz.shape[1]
    out = np.empty((b, m, nsample), np.int32)
    lib().oracle_ball_query(_p(new_xyz), _p(xyz), _ci(b), _ci(n), _ci(m), _cf(radius),
                            _ci(nsample), _p(out))
    return out


def cylinder_query(new_xyz, xyz, rot, radius, hmin, hmax, nsample):
    """cylinder_query_gpu.cu:9-57 (pybind order, cylinder_query.cpp:10)."""
    new_xyz, xyz, rot = _f32(new_xyz), _f32(xyz), _f32(rot)
    b, m, _ = new_xyz.shape
    n = xyz.shape[1]
    out = np.empty((b, m, nsample), np.int32)
    lib().oracle_cylinder_query(_p(new_xyz), _p(xyz), _p(rot), _ci(b), _ci(n), _ci(m),
                                _cf(radius), _cf(hmin), _cf(hmax), _ci(nsample), _p(out))
    return out


def group_points(points, idx):
    """group_points_gpu.cu:13-33. points (B,C,N), idx (B,M,ns) -> (B,C,M,ns)."""
    points, idx = _f32(points), _i32(idx)
    b, c, n = points.shape
    _, m, ns = idx.shape
    out = np.empty((b, c, m, ns), np.float32)
    lib().oracle_group_points(_p(points), _p(idx), _ci(b), _ci(c), _ci(n), _ci(m), _ci(ns), _p(out))
    return out


def group_points_grad(grad_out, idx, n):
    """group_points_gpu.cu:48-69. grad_out (B,C,M,ns) -> (B,C,n)."""
    grad_out, idx = _f32(grad_out), _i32(idx)
    b, c, m, ns = grad_out.shape
    out = np.empty((b, c, n), np.float32)
    lib().oracle_group_points_grad(_p(grad_out), _p(idx), _ci(b), _ci(c), _ci(n), _ci(m), _ci(ns),
                                   _p(out))
    return out


def three_nn(unknown, known):
    """interpolate_gpu.cu:14-64 -> (dist2 (B,n,3) f32, idx (B,n,3) i32); NO sqrt (that is
    pointnet2_utils.py:140, done by the host wrapper)."""
    unknown, known = _f32(unknown), _f32(known)
    b, n, _ = unknown.shape
    m = known.shape[1]
    d = np.empty((b, n, 3), np.float32)
    i = np.empty((b, n, 3), np.int32)
    lib().oracle_three_nn(_p(unknown), _p(known), _ci(b), _ci(n), _ci(m), _p(d), _p(i))
    return d, i


def three_interpolate(points, idx, weight):
    """interpolate_gpu.cu:77-106. points (B,c,m), idx/weight (B,n,3) -> (B,c,n)."""
    points, idx, weight = _f32(points), _i32(idx), _f32(weight)
    b, c, m = points.shape
    n = idx.shape[1]
    out = np.empty((b, c, n), np.float32)
    lib().oracle_three_interpolate(_p(points), _p(idx), _p(weight), _ci(b), _ci(c), _ci(m), _ci(n),
                                   _p(out))
    return out


def three_interpolate_grad(grad_out, idx, weight, m):
    """interpolate_gpu.cu:121-148. grad_out (B,c,n) -> (B,c,m)."""
    grad_out, idx, weight = _f32(grad_out), _i32(idx), _f32(weight)
    b, c, n = grad_out.shape
    out = np.empty((b, c, m), np.float32)
    lib().oracle_three_interpolate_grad(_p(grad_out), _p(idx), _p(weight), _ci(b), _ci(c), _ci(n),
                                        _ci(m), _p(out))
    return out


def knn(ref, query, k=1, fma_chain=True):
    """knn/src/cuda/knn.cu:36-176 (fma_chain) or knn/src/cpu/knn_cpu.cpp:4-56 (not).
    ref (B,D,Nr), query (B,D,Nq) -> (B,k,Nq) int64, 1-based (knn_modules.py:9-17)."""
    ref, query = _f32(ref), _f32(query)
    b, d, nr = ref.shape
    nq = query.shape[2]
    out = np.zeros((b, k, nq), np.int64)
    lib().oracle_knn(_p(ref), _p(query), _ci(b), _ci(d), _ci(nr), _ci(nq), _ci(k),
                     _ci(1 if fma_chain else 0), _p(out))
    return out
