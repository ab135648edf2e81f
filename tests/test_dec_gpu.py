"""GPU parity of the fused DecGridRL kernel: golden traces from the live reference, replayed
through the C ABI (VecDecGridRL -> ctypes -> libmcov_b200.so), plus randomized cases against the
oracle.  Bit-exact: positions, planes, done, coverage, float32 observations, float64 rewards."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, golden_names
from oracle.trace import Trace

pytestmark = pytest.mark.gpu


DEC_TRACES = [n for n in golden_names("dec")]


@pytest.mark.parametrize("name", DEC_TRACES)
def test_dec_golden_single_env(name):
    from marl_coverage_b200 import VecDecGridRL
    from gpu_helpers import replay_trace_on_gpu
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    tol = 1e-6 if tr.config["mini_map_rad"] > 0 else 0.0
    replay_trace_on_gpu(tr, VecDecGridRL, num_envs=1, check_envs=(0,), obs_tol=tol)


@pytest.mark.parametrize("name", ["dec_c2_square_100", "dec_lidar_21x10_100", "dec_square_thresh_incr",
                                  "dec_mapsharing_comm12"])
def test_dec_golden_batched_copies(name):
    """The same session in 67 envs at once: catches cross-env indexing / persistent-loop bugs."""
    from marl_coverage_b200 import VecDecGridRL
    from gpu_helpers import replay_trace_on_gpu
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    replay_trace_on_gpu(tr, VecDecGridRL, num_envs=67, check_envs=(0, 31, 66))


@pytest.mark.parametrize("group", [4, 8, 16, 32])
@pytest.mark.parametrize("name", ["dec_c2_square_100", "dec_lidar_21x10_100", "dec_mapsharing_comm12",
                                  "dec_square_thresh_incr", "dec_single_square_tool"])
def test_dec_golden_every_group_width(name, group, monkeypatch):
    """The kernel picks 4/8/16/32 lanes per env from the batch size; every width must give the same
    bits (MCOV_DEC_GROUP pins it)."""
    from marl_coverage_b200 import VecDecGridRL
    from gpu_helpers import replay_trace_on_gpu
    monkeypatch.setenv("MCOV_DEC_GROUP", str(group))
    monkeypatch.setenv("MCOV_DEC_SHAPE", "0")
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    replay_trace_on_gpu(tr, VecDecGridRL, num_envs=37, check_envs=(0, 36))


@pytest.mark.parametrize("name,group,store", [(n, g, 1) for n in ("dec_c2_square_100", "dec_lidar_21x10_100",
                                                                   "dec_mapsharing_comm12", "dec_dist_reward_15")
                                              for g in (8, 16, 32)] + [("dec_c2_square_100", 8, 0)])
def test_dec_golden_store_warp(name, group, store, monkeypatch):
    """Large batches hand their observation blocks to a store warp (bit-stream slots, mbarriers, TMA bulk
    stores); small ones store directly.  Both paths must give the same bits (MCOV_DEC_STORE pins the path).
    Two CTAs loop over 150 envs, so every lane group reuses both of its slots several times (the one
    store = 0 case runs the same capped grid through the direct stores)."""
    from marl_coverage_b200 import VecDecGridRL
    from gpu_helpers import replay_trace_on_gpu
    monkeypatch.setenv("MCOV_DEC_SHAPE", "0")
    monkeypatch.setenv("MCOV_DEC_GROUP", str(group))
    monkeypatch.setenv("MCOV_DEC_STORE", str(store))
    monkeypatch.setenv("MCOV_DEC_GRID", "2")
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    replay_trace_on_gpu(tr, VecDecGridRL, num_envs=150, check_envs=(0, 61, 149))


@pytest.mark.parametrize("group,team", [(8, 16), (8, 32), (16, 32)])
@pytest.mark.parametrize("name", ["dec_c2_square_100", "dec_lidar_21x10_100", "dec_mapsharing_comm12",
                                  "dec_dist_reward_15"])
def test_dec_golden_store_teams(name, group, team, monkeypatch):
    """Two or four lane groups store their observation blocks together, 16 / 32 lanes per block (MCOV_DEC_TEAM pins
    the team; the kernel picks it from the shape).  Two CTAs loop over 150 envs: the last teams are incomplete."""
    from marl_coverage_b200 import VecDecGridRL
    from gpu_helpers import replay_trace_on_gpu
    monkeypatch.setenv("MCOV_DEC_GROUP", str(group))
    monkeypatch.setenv("MCOV_DEC_TEAM", str(team))
    monkeypatch.setenv("MCOV_DEC_GRID", "2")
    monkeypatch.setenv("MCOV_DEC_SHAPE", "0")
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    replay_trace_on_gpu(tr, VecDecGridRL, num_envs=150, check_envs=(0, 61, 149))


@pytest.mark.parametrize("group", [8, 16, 32])
@pytest.mark.parametrize("name", [n for n in DEC_TRACES if "lidar" not in n])
def test_dec_golden_map_table_in_shared_memory(name, group, monkeypatch):
    """Square-sensor handles can keep the whole read-only truth-map table in shared memory (one 1024-thread CTA per SM
    stages it once per launch with TMA bulk copies; MCOV_DEC_SMT pins it).  Every non-lidar golden trace must give the
    same bits; two CTAs loop over 150 envs."""
    from marl_coverage_b200 import VecDecGridRL
    from gpu_helpers import replay_trace_on_gpu
    monkeypatch.setenv("MCOV_DEC_SHAPE", "0")
    monkeypatch.setenv("MCOV_DEC_GROUP", str(group))
    monkeypatch.setenv("MCOV_DEC_SMT", "1")
    monkeypatch.setenv("MCOV_DEC_GRID", "2")
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    tol = 1e-6 if tr.config["mini_map_rad"] > 0 else 0.0
    replay_trace_on_gpu(tr, VecDecGridRL, num_envs=150, check_envs=(0, 61, 149), obs_tol=tol)


@pytest.mark.parametrize("tma", [0, 1])
@pytest.mark.parametrize("name", DEC_TRACES)
def test_dec_golden_cta_per_env_shape(name, tma, monkeypatch):
    """Small batches / large maps run one CTA per env (warp per robot, lane per band row; mcov_dec_cta.cu), with
    the map windows staged into shared memory by bulk async copies (tma = 1) or read with plain loads (tma = 0).
    Every golden trace must give the same bits as through the lane-group kernel (MCOV_DEC_SHAPE pins the shape).
    Three CTAs loop over seven envs."""
    from marl_coverage_b200 import VecDecGridRL
    from gpu_helpers import replay_trace_on_gpu
    monkeypatch.setenv("MCOV_DEC_SHAPE", "1")
    monkeypatch.setenv("MCOV_DEC_TMA", str(tma))
    monkeypatch.setenv("MCOV_DEC_GRID", "3")
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    tol = 1e-6 if tr.config["mini_map_rad"] > 0 else 0.0
    replay_trace_on_gpu(tr, VecDecGridRL, num_envs=7, check_envs=(0, 6), obs_tol=tol)


@pytest.mark.parametrize("name", [n for n in DEC_TRACES if "lidar" in n])
def test_dec_lidar_float64_march_path(name, monkeypatch):
    """The lidar normally walks an integer beam table proved exact on the host; the float64 DDA
    march it replaces stays as the fallback and must give the same bits (MCOV_LIDAR_FP64=1)."""
    from marl_coverage_b200 import VecDecGridRL
    from gpu_helpers import replay_trace_on_gpu
    monkeypatch.setenv("MCOV_LIDAR_FP64", "1")
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    replay_trace_on_gpu(tr, VecDecGridRL, num_envs=3, check_envs=(0, 2))

