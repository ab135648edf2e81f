"""GPU parity of the fused SuperGridRL kernel: golden traces from the live reference replayed
through the C ABI (VecSuperGridRL -> ctypes -> libmcov_b200.so).  Bit-exact: positions, covered /
observed-obstacle planes, done, coverage, thresholds, the full float32 state tensor (CRC per step,
full tensors at checkpoints, incl. the L1 distance channel) and float64 rewards — also with the
fractional free_penalty = 0.2, because the kernel adds in the reference's order."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, golden_names
from oracle.trace import Trace

pytestmark = pytest.mark.gpu

SUPER_TRACES = golden_names("super")


@pytest.mark.parametrize("name", SUPER_TRACES)
def test_super_golden_single_env(name):
    from marl_coverage_b200 import VecSuperGridRL
    from gpu_helpers import replay_trace_on_gpu
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    replay_trace_on_gpu(tr, VecSuperGridRL, num_envs=1, check_envs=(0,))


@pytest.mark.parametrize("name", ["super_c1_100", "super_scanning_n5", "super_dist_reward_15"])
def test_super_golden_batched_copies(name):
    """The same session in 301 envs (> one wave of CTAs): persistent-loop / indexing bugs."""
    from marl_coverage_b200 import VecSuperGridRL
    from gpu_helpers import replay_trace_on_gpu
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    replay_trace_on_gpu(tr, VecSuperGridRL, num_envs=701, check_envs=(0, 300, 700))
