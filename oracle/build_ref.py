#!/usr/bin/env python
"""Recipe that compiles the reference's LEGACY CUDA extensions where they lie.

TEST INFRASTRUCTURE ONLY.  Nothing here is product code and nothing here is
copied into the repo: the sources are compiled in place from /root/reference
(read-only) and only the binaries land in oracle/_ref/ (git-ignored, but NOT
gpurun-ignored, so they travel to the GPU box).

  oracle/_ref/pointnet2/_ext*.so        <- /root/reference/pointnet2/_ext_src/src/*.{cpp,cu}
  oracle/_ref/knn_pytorch/knn_pytorch*.so <- /root/reference/knn/src/{vision.cpp,cpu/knn_cpu.cpp,cuda/knn.cu}

The reference's own build system (setup.py / BuildExtension) is NOT run; this
script calls g++/nvcc directly with the flags pointnet2/setup.py:21-28 and
knn/setup.py:28-39 would pass (-O2, WITH_CUDA, default --fmad), targeting
sm_100 (plain, as torch's arch list would; the legacy kernels use no
arch-specific feature).  On the GPU box these two modules are
  * the primary parity oracle for every pointnet2 op and for knn (bit-exact
    index comparisons against the very kernels this repo replaces), and
  * the "legacy CUDA ops on the same B200" timed baseline of BASELINE.md §3.
If /root/reference is absent (the GPU box) this script is a no-op.
"""
import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")
OBJ = os.path.join(OUT, "obj")


def _flags():
    import torch
    from torch.utils import cpp_extension as ce
    inc = ce.include_paths("cuda") + [sysconfig.get_paths()["include"]]
    common = [f"-I{p}" for p in inc] + [
        "-DTORCH_API_INCLUDE_EXTENSION_H",
        f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
        "-O2", "-std=c++17",
    ]
    libdirs = ce.library_paths("cuda")
    link = [f"-L{p}" for p in libdirs] + [f"-Wl,-rpath,{p}" for p in libdirs] + [
        "-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda", "-lcudart"]
    return common, link


def _compile(src, obj, extra, common):
    if os.path.exists(obj) and os.path.getmtime(obj) >= os.path.getmtime(src):
        return
    if src.endswith(".cu"):
        cmd = ["nvcc", "-c", src, "-o", obj, "-gencode", "arch=compute_100,code=sm_100",
               "-Xcompiler", "-fPIC", "-w",
               "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__",
               "-D__CUDA_NO_HALF2_OPERATORS__"] + common + extra
    else:
        cmd = ["g++", "-c", src, "-o", obj, "-fPIC", "-w"] + common + extra
    subprocess.check_call(cmd)


def build(verbose=True):
    if not os.path.isdir(REF):
        if verbose:
            print(f"[oracle/_ref] {REF} absent: using prebuilt files (if any)")
        return False
    os.makedirs(OBJ, exist_ok=True)
    common, link = _flags()
    suffix = sysconfig.get_config_var("EXT_SUFFIX")
    jobs = []
    # --- pointnet2._ext -------------------------------------------------
    p2 = os.path.join(REF, "pointnet2", "_ext_src")
    p2_srcs = sorted(os.path.join(p2, "src", f) for f in os.listdir(os.path.join(p2, "src"))
                     if f.endswith((".cpp", ".cu")))
    p2_extra = [f"-I{os.path.join(p2, 'include')}", "-DTORCH_EXTENSION_NAME=_ext"]
    p2_objs = []
    for s in p2_srcs:
        o = os.path.join(OBJ, "p2_" + os.path.basename(s).replace(".", "_") + ".o")
        p2_objs.append(o)
        jobs.append((s, o, p2_extra))
    # --- knn_pytorch.knn_pytorch ---------------------------------------
    kn = os.path.join(REF, "knn", "src")
    kn_srcs = [os.path.join(kn, "vision.cpp"), os.path.join(kn, "cpu", "knn_cpu.cpp"),
               os.path.join(kn, "cuda", "knn.cu")]
    kn_extra = [f"-I{kn}", "-DTORCH_EXTENSION_NAME=knn_pytorch", "-DWITH_CUDA", "-DCUDA_HAS_FP16=1"]
    kn_objs = []
    for s in kn_srcs:
        o = os.path.join(OBJ, "knn_" + os.path.basename(s).replace(".", "_") + ".o")
        kn_objs.append(o)
        jobs.append((s, o, kn_extra))
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(lambda j: _compile(j[0], j[1], j[2], common), jobs))
    for pkg, name, objs in (("pointnet2", "_ext", p2_objs), ("knn_pytorch", "knn_pytorch", kn_objs)):
        d = os.path.join(OUT, pkg)
        os.makedirs(d, exist_ok=True)
        open(os.path.join(d, "__init__.py"), "a").close()
        so = os.path.join(d, name + suffix)
        if not os.path.exists(so) or any(os.path.getmtime(o) > os.path.getmtime(so) for o in objs):
            subprocess.check_call(["g++", "-shared", "-o", so] + objs + link)
        if verbose:
            print("[oracle/_ref] built", so)
    return True


if __name__ == "__main__":
    sys.exit(0 if build() else 0)
