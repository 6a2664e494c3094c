"""N>1 host logic on CPU: world_size-2 gloo run of the scene sharding + whole-job reduction bench.py uses."""
import importlib.util
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_sharding():
    # loaded by path: importing the package itself needs the CUDA library
    spec = importlib.util.spec_from_file_location("b200_sharding", os.path.join(ROOT, "graspnet-baseline_b200", "sharding.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = load_sharding()
    mine = sh.scene_shard(7, rank, world)
    total, ms, sps = sh.whole_job_throughput(len(mine), 10.0 * (rank + 1))
    q.put((rank, mine, total, ms, sps))
    dist.destroy_process_group()


def test_two_rank_sharding_and_reduction():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, s0, t0, m0, v0), (r1, s1, t1, m1, v1) = res
    assert s0 == [0, 2, 4, 6] and s1 == [1, 3, 5]
    assert sorted(s0 + s1) == list(range(7))                 # every scene exactly once, nothing exchanged
    assert t0 == t1 == 7.0 and m0 == m1 == 20.0               # SUM of scenes, MAX of time on every rank
    assert abs(v0 - 7.0 / 0.020) < 1e-9 and v0 == v1


def test_single_process_is_identity():
    sh = load_sharding()
    assert sh.scene_shard(5, 0, 1) == [0, 1, 2, 3, 4]
    assert sh.whole_job_throughput(4, 8.0) == (4.0, 8.0, 500.0)
