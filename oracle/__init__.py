"""CPU parity oracle for the point-cloud geometry hot path.

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs -- never from the product
package (graspnet-baseline_b200/), which must fail loudly without its CUDA
library instead of falling back to anything in here.
"""
from .cpu import *  # noqa: F401,F403
from .post import *  # noqa: F401,F403
from . import cpu, post, mlp  # noqa: F401
