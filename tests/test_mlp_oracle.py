"""The fused group -> shared MLP -> max-pool oracle (oracle/mlp.py) against the outputs of the reference's own
SharedMLP (tests/golden/mlp_*.npz, made by tests/golden/make_golden.py in the build container).  CPU only."""
import glob
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
RTOL = 1e-5      # north_star: features within 1e-5 relative (fp32)


def load_case(path, b200_synth, oracle_mod):
    z = np.load(path)
    params = b200_synth.mlp_params(int(z["seed"]), [int(d) for d in z["dims"]])
    layers = [oracle_mod.mlp.fold_bn(*p) for p in params]
    rot = z["rot"] if z["rot"].size else None
    feat = z["feat"] if z["feat"].size else None
    return z, layers, rot, feat


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(HERE, "golden", "mlp_*.npz"))))
def test_oracle_matches_reference_shared_mlp(path, oracle_mod):
    import importlib.util
    spec = importlib.util.spec_from_file_location("b200_synth", os.path.join(os.path.dirname(HERE), "graspnet-baseline_b200", "synth.py"))
    synth = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(synth)
    z, layers, rot, feat = load_case(path, synth, oracle_mod)
    got = oracle_mod.mlp.group_mlp_maxpool(z["xyz"], z["new_xyz"], z["idx"], layers, float(z["radius"]), rot, feat)
    want = z["out"]
    assert got.shape == want.shape
    assert np.allclose(got, want, rtol=RTOL, atol=RTOL * np.abs(want).max())


def test_golden_files_present():
    assert len(glob.glob(os.path.join(HERE, "golden", "mlp_*.npz"))) == 5
