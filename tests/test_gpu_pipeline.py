"""The whole-scene pass: serial, multi-stream and CUDA-graph orchestration give identical results,
and a replayed graph follows new inputs copied into its static buffers."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def flat(out):
    r = {}
    for k, v in out.items():
        if isinstance(v, list):
            for i, e in enumerate(v):
                if isinstance(e, tuple):
                    keep, nkeep = e
                    r["%s_%d" % (k, i)] = keep[: int(nkeep.item())].cpu().numpy()
                else:
                    r["%s_%d" % (k, i)] = e.cpu().numpy()
        else:
            r[k] = v.cpu().numpy()
    return r


@pytest.mark.parametrize("batch", [1, 2])
def test_graph_equals_serial_and_follows_inputs(b200, batch):
    from b200grasp.pipeline import GeometryScene
    cfg = [(256, 0.06, 32, 0), (128, 0.12, 16, 16), (64, 0.2, 16, 32), (32, 0.3, 16, 32)]
    sc = GeometryScene(batch=batch, num_point=3000, num_grasp=200, seed=7, device="cuda", sa_cfg=cfg)
    want = flat(sc.run())
    got_s = flat(sc.run_streams())
    torch.cuda.synchronize()
    got_g = flat(sc.run_graph())
    assert set(want) == set(got_s) == set(got_g)
    for k in want:
        assert np.array_equal(want[k], got_s[k]), k
        assert np.array_equal(want[k], got_g[k]), k
    assert sc.launches > 0
    # new scene through the same graph
    other = GeometryScene(batch=batch, num_point=3000, num_grasp=200, seed=99, device="cuda", sa_cfg=cfg)
    sc.cloud.copy_(other.cloud)
    sc.rot.copy_(other.rot)
    for k in sc.feat:
        sc.feat[k].copy_(other.feat[k])
    for a, b in zip(sc.grasps, other.grasps):
        a.copy_(b)
    for a, b in zip(sc.ngrasp, other.ngrasp):
        a.copy_(b)
    want2 = flat(other.run())
    got2 = flat(sc.run_graph())
    for k in want2:
        assert np.array_equal(want2[k], got2[k]), k


def test_streams_with_duplicate_points_prefix_checks_fail(b200, oracle_mod):
    """A cloud of 100 distinct points, each 30 times: SA1's FPS returns 100 distinct samples and then repeats, so
    SA2's input is no FPS-ordered prefix and SA3/SA4 (whose own checks could pass) must not take the shortcut on
    the strength of their own verdict alone.  Serial (checks inside each FPS call), streams and graph (checks
    launched side by side after SA1) and the oracle all agree."""
    from b200grasp.pipeline import GeometryScene
    cfg = [(256, 0.06, 32, 0), (128, 0.12, 16, 16), (64, 0.2, 16, 32), (32, 0.3, 16, 32)]
    sc = GeometryScene(batch=2, num_point=3000, num_grasp=100, seed=3, device="cuda", sa_cfg=cfg)
    cloud = sc.host["cloud"].copy()
    cloud[1] = np.tile(cloud[1][:100], (30, 1))
    sc.cloud.copy_(torch.from_numpy(cloud))
    want = flat(sc.run())
    got_s = flat(sc.run_streams())
    torch.cuda.synchronize()
    got_g = flat(sc.run_graph())
    for k in want:
        assert np.array_equal(want[k], got_s[k]), k
        assert np.array_equal(want[k], got_g[k]), k
    cur = cloud
    for lvl, (npoint, _, _, _) in enumerate(cfg):
        inds = oracle_mod.furthest_point_sampling(cur, npoint)
        assert np.array_equal(want["sa%d_inds" % (lvl + 1)], inds), lvl
        cur = np.stack([cur[i][inds[i]] for i in range(2)])
    assert not np.array_equal(want["sa2_inds"][1], np.arange(128))
    assert np.array_equal(want["sa2_inds"][0], np.arange(128))
