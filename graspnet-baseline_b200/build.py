#!/usr/bin/env python
"""Builds libb200grasp.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python graspnet-baseline_b200/build.py [--force] [--verbose]

One translation unit per kernel family; objects are cached under csrc/build/ and
rebuilt when the source (or a header) is newer.  No torch headers are involved:
the library is plain CUDA runtime + extern "C".
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
LIB = os.path.join(HERE, "lib", "libb200grasp.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC"]


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")]
    hdrs.append(os.path.join(os.path.dirname(HERE), "include", "b200grasp.h"))
    hdr_m = max(os.path.getmtime(h) for h in hdrs)
    jobs, objs = [], []
    for f in sources():
        src = os.path.join(CSRC, f)
        obj = os.path.join(OBJ, f[:-3] + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_m):
            jobs.append([NVCC, "-c", src, "-o", obj] + FLAGS + (["-Xptxas", "-v"] if verbose else []))

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode:
            sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode:
            raise RuntimeError("nvcc failed: " + cmd[2])

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(run, jobs))
    if jobs or not os.path.exists(LIB):
        run([NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    return LIB


def build_fps_profile():
    """Instrumented FPS-only library (clock64 stage counters) for tools/fps_prof.py; not shipped."""
    out = os.path.join(HERE, "lib", "libb200fps_prof.so")
    cmd = [NVCC, "-shared", "-o", out, os.path.join(CSRC, "fps.cu"), os.path.join(CSRC, "fps_multi.cu"), os.path.join(CSRC, "common.cu"),
           "-DFPS_PROFILE"] + FLAGS
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed (fps profile build)")
    return out


def build_mlp_profile():
    """Instrumented fused-MLP library (clock64 timeline of CTA 0) for tools/mlp_prof.py; not shipped."""
    out = os.path.join(HERE, "lib", "libb200mlp_prof.so")
    cmd = [NVCC, "-shared", "-o", out, os.path.join(CSRC, "mlp_fused.cu"), os.path.join(CSRC, "mlp_f16.cu"),
           os.path.join(CSRC, "common.cu"), "-DMLP_PROFILE"] + FLAGS + [a for a in sys.argv if a.startswith("-D")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed (mlp profile build)")
    return out


if __name__ == "__main__":
    if "--mlp-profile" in sys.argv:
        print(build_mlp_profile())
    elif "--fps-profile" in sys.argv:
        print(build_fps_profile())
    else:
        print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
