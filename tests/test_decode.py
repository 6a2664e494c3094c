"""pred_decode (models/graspnet.py:77-134): the oracle against golden vectors produced by the reference's own
function (CPU, tests/golden/make_golden.py), and the CUDA kernel against the oracle / the golden vectors.
Selected columns (score, width, height, depth, centre, object id) are bit-exact; the rotation goes through norms,
sin/cos and a 3x3 product: 1e-6 absolute (entries are <= 1 in magnitude)."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KEYS = ("objectness_score", "grasp_score_pred", "grasp_angle_cls_pred", "grasp_width_pred", "grasp_tolerance_pred",
        "fp2_xyz", "grasp_top_view_xyz")
EXACT = [0, 1, 2, 3, 13, 14, 15, 16]
ROT_ATOL = 1e-6


def check_rows(got, want):
    assert got.shape == want.shape and got.dtype == np.float32
    assert (got[:, EXACT] == want[:, EXACT]).all()
    assert np.abs(got[:, 4:13] - want[:, 4:13]).max(initial=0.0) <= ROT_ATOL


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_oracle_pred_decode_equals_reference_golden(oracle_mod, name):
    z = np.load(os.path.join(GOLD, "decode_%s.npz" % name))
    for i in range(z["objectness_score"].shape[0]):
        check_rows(oracle_mod.post.pred_decode(*[z[k][i] for k in KEYS]), z["rows_%d" % i])


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_pred_decode_kernel_equals_reference_golden(b200, name):
    import torch
    z = np.load(os.path.join(GOLD, "decode_%s.npz" % name))
    ep = {k: torch.from_numpy(z[k]).cuda() for k in KEYS}
    out = b200.pred_decode(ep)
    assert len(out) == z["objectness_score"].shape[0]
    for i, o in enumerate(out):
        check_rows(o.cpu().numpy(), z["rows_%d" % i])
    rows, nrows = b200.pred_decode_device(ep)
    assert nrows.cpu().tolist() == [len(o) for o in out]


@pytest.mark.gpu
@pytest.mark.parametrize("b,ns,a,d,seed", [(1, 1024, 12, 4, 0), (3, 3000, 12, 4, 1), (2, 77, 5, 3, 2), (1, 1, 12, 4, 3),
                                          (1, 2048, 12, 8, 4)])
def test_pred_decode_kernel_equals_oracle(b200, oracle_mod, b, ns, a, d, seed):
    """sizes beyond one 1024-seed pass, other head shapes, every-seed / no-seed objectness."""
    import torch
    rng = np.random.default_rng(seed)
    f = np.float32
    obj = rng.standard_normal((b, 2, ns)).astype(f)
    if seed == 1:
        obj[0, 1] = obj[0, 0] + 1      # scene 0: every seed is an object
        obj[1, 1] = obj[1, 0]          # scene 1: ties -> argmax 0 -> none
    heads = [obj, (rng.random((b, a, ns, d)) * 1.2 - 0.1).astype(f), np.round(rng.standard_normal((b, a, ns, d)) * 2).astype(f),
             (rng.random((b, a, ns, d)) * 0.12 - 0.01).astype(f), (rng.random((b, a, ns, d)) * 0.06).astype(f),
             (rng.random((b, ns, 3)) * 0.6 - 0.3).astype(f)]
    v = rng.standard_normal((b, ns, 3)).astype(f)
    v[:, ::9] = [0, 0, 1]
    heads.append(v)
    ep = {k: torch.from_numpy(h).cuda() for k, h in zip(KEYS, heads)}
    out = b200.pred_decode(ep)
    for i in range(b):
        check_rows(out[i].cpu().numpy(), oracle_mod.post.pred_decode(*[h[i] for h in heads]))


@pytest.mark.gpu
@pytest.mark.parametrize("ns,kept", [(1024, (5, 300, 900)), (1024, (0, 1, 2)), (200, (127, 128, 199)), (3, (0, 1, 2))])
def test_pred_decode_three_surviving_seeds(b200, oracle_mod, ns, kept):
    """exactly three kept seeds: the reference's torch.cross without `dim` runs down the rows (loss_utils.py:89);
    the three seeds lie in different CTAs of the decode kernel"""
    import torch
    rng = np.random.default_rng(ns + kept[1])
    f = np.float32
    a, d = 12, 4
    obj = np.zeros((1, 2, ns), f)
    obj[0, 0] = 1
    obj[0, 1, list(kept)] = 2
    heads = [obj, (rng.random((1, a, ns, d)) * 1.2 - 0.1).astype(f), rng.standard_normal((1, a, ns, d)).astype(f),
             (rng.random((1, a, ns, d)) * 0.12 - 0.01).astype(f), (rng.random((1, a, ns, d)) * 0.06).astype(f),
             (rng.random((1, ns, 3)) * 0.6 - 0.3).astype(f), rng.standard_normal((1, ns, 3)).astype(f)]
    ep = {k: torch.from_numpy(h).cuda() for k, h in zip(KEYS, heads)}
    out = b200.pred_decode(ep)
    want = oracle_mod.post.pred_decode(*[h[0] for h in heads])
    assert want.shape == (3, 17)
    check_rows(out[0].cpu().numpy(), want)
