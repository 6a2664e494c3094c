import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_sessionstart(session):
    """B200_STACK_ARITH=fp32 runs the suite with the exact-fp32 FP / ApproachNet stacks (fused_modules.STACK_ARITH)"""
    if os.environ.get("B200_STACK_ARITH"):
        import b200grasp
        b200grasp.fused_modules.STACK_ARITH = os.environ["B200_STACK_ARITH"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle
    oracle.cpu.build()
    return oracle


@pytest.fixture(scope="session")
def b200():
    import b200grasp
    return b200grasp


@pytest.fixture(scope="session")
def legacy():
    """The reference's own legacy CUDA extensions built by oracle/build_ref.py (None if absent)."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "pointnet2")):
        return None
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    if ref_dir not in sys.path:
        sys.path.insert(0, ref_dir)
    try:
        import pointnet2._ext as ext
        from knn_pytorch import knn_pytorch
    except Exception:
        return None
    return {"ext": ext, "knn": knn_pytorch}
