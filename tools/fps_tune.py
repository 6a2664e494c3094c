#!/usr/bin/env python
"""Sweep the FPS cluster size / CTA size on the GPU box and check every variant against the oracle.

    python tools/fps_tune.py            # prints one line per (shape, C, NT): us per call, ns per step
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200grasp  # noqa: E402
import oracle  # noqa: E402

lib = b200grasp._abi.lib


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3


def main():
    oracle.cpu.build()
    shapes = [(1, 20000, 2048), (1, 2048, 1024), (1, 1024, 512), (1, 512, 256), (64, 20000, 2048), (1, 200000, 4096)]
    for b, n, m in shapes:
        x = np.stack([b200grasp.synth.adversarial_cloud(n, 100 + i) for i in range(b)])
        want = oracle.furthest_point_sampling(x[:2], m) if n <= 20000 else None
        xd = torch.from_numpy(x).cuda()
        for nt in (128, 256, 512):
            for c in (0, 1, 2, 4, 8, 16):
                lib.b200_fps_set_cluster(c)
                lib.b200_fps_set_threads(nt)
                try:
                    got = b200grasp.ops.furthest_point_sampling(xd, m)
                    torch.cuda.synchronize()
                except RuntimeError as e:
                    print("b=%d n=%d m=%d C=%d NT=%d: %s" % (b, n, m, c, nt, str(e)[:60]))
                    continue
                ok = "" if want is None else ("ok" if (got[:2].cpu().numpy() == want).all() else "MISMATCH")
                us = timeit(lambda: b200grasp.ops.furthest_point_sampling(xd, m))
                print("b=%d n=%d m=%d C=%d NT=%d: %9.1f us  %7.1f ns/step %s" % (b, n, m, c, nt, us, us * 1e3 / (m - 1), ok),
                      flush=True)
    lib.b200_fps_set_cluster(0)
    lib.b200_fps_set_threads(0)


if __name__ == "__main__":
    main()
