"""Host-side logic that needs no GPU: map supply, plane packing, action forms, lidar beam table,
algorithmic-byte model of bench.py."""
import numpy as np
import pytest

from marl_coverage_b200 import maps
from marl_coverage_b200.action_spaces import Discrete
from marl_coverage_b200.envs import _scalar_action
from marl_coverage_b200.sensors import LidarSensor, SquareSensor, lidar_beam_table
from oracle import coverage_oracle as oc


def test_plane_pack_roundtrip():
    rng = np.random.default_rng(0)
    for W, L in ((1, 1), (5, 31), (7, 32), (9, 33), (102, 102), (64, 65)):
        m = rng.random((W, L)) < 0.4
        w = maps.pack_plane_host(m)
        strips, Wp, PW = maps._geometry(W, L)
        assert w.shape == (PW,) and Wp % 4 == 0
        assert np.array_equal(maps.unpack_plane_host(w, W, L), m)
        # bit b of word (s, x) is cell (x, 32 s + b)
        x, y = W - 1, L - 1
        assert bool((w[(y >> 5) * Wp + x] >> (y & 31)) & 1) == bool(m[x, y])


def test_builtin_and_committed_grid_sets():
    train, test = maps.gridload(None)
    assert len(train) == 6 and len(test) == 3 and all(g.shape == (15, 15) for g in train + test)
    g100 = maps.load_grid_set("bg2_100x100")
    assert len(g100) == 75 and g100[0].shape == (100, 100)
    assert set(np.unique(g100[0])) <= {-1.0, 1.0}
    tr, te = maps.gridload({"grid_dir": "./Grids/bg2_100x100", "numgrids": 20})
    assert len(tr) == 10 and len(te) == 10             # gridmaker.py:96-102 half split
    cells, values = maps.grids_to_cells([np.full((4, 4), 0.5), np.array([[-2.0, 0.0, 3.0, 1.0]] * 4)])
    assert cells[0].tolist() == [[1] * 4] * 4 and cells[1][0].tolist() == [-1, 0, 1, 1]
    assert values[0, 0, 0] == 0.5 and values[1][0].tolist() == [0.0, 0.0, 3.0, 1.0]      # free cells keep their worth
    assert maps.grids_to_cells(g100[:2])[1] is None                                       # +-1 PNG maps need no table
    with pytest.raises(ValueError):
        maps.grids_to_cells([np.full((4, 4), np.nan)])


def test_gridgen_matches_reference_contract():
    np.random.seed(3)
    tr, te = maps.gridgen({"prob_obst": 0.2, "gridwidth": 12, "gridlen": 9, "numgrids": 4})
    assert len(tr) == 4 and all(g.shape == (12, 9) for g in tr)
    assert set(np.unique(np.stack(tr))) <= {-1.0, 1.0}


def test_action_forms():
    import torch
    assert _scalar_action(None) == ("sentinel", -1)
    assert _scalar_action(-1) == ("sentinel", -1)
    assert _scalar_action(np.int64(27)) == ("joint", 27)
    assert _scalar_action(torch.tensor(5)) == ("joint", 5)
    kind, v = _scalar_action([0, 3, 2])
    assert kind == "per_robot" and v.tolist() == [0, 3, 2]
    np.random.seed(0)
    d = Discrete(256)
    s = d.sample()
    assert 0 <= int(s) < 256 and d.sample(5).shape == (5,)
    assert oc.decode_base4(27, 4).tolist() == [3, 2, 1, 0]


def test_lidar_beam_table_is_the_oracles():
    for n in (1, 5, 21, 31):
        assert np.array_equal(lidar_beam_table(n), oc.lidar_beam_table(n))
    with pytest.raises(AssertionError):
        LidarSensor({"num_lasers": 20, "range": 4})
    assert SquareSensor({"range": 2}).env_keys()["sensor_config"] == {"range": 2}


def test_algorithmic_bytes_match_survey():
    import bench
    kind, grid, E, cfg, _ = bench.WORKLOADS["c2"]
    assert bench.alg_bytes_per_env_step(kind, cfg, 102, 102) == 7769
    kind, grid, E, cfg, _ = bench.WORKLOADS["c3"]
    assert bench.alg_bytes_per_env_step(kind, cfg, 258, 258) == 23341


def test_reference_arm_prints_the_contract_line():
    """bench.py --impl reference runs on the host cores only: the JSON line must carry every key the driver
    reads (metric / unit of BASELINE.json, cpu_baseline, e2e with zero copies, no GPU launches)."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    import time
    t0 = time.perf_counter()
    res = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "3",
                          "--warmup", "1", "--sample-seconds", "1.0"], capture_output=True, text=True, timeout=300,
                         cwd=root)
    wall = time.perf_counter() - t0
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    with open(os.path.join(root, "BASELINE.json")) as f:
        base = json.load(f)
    assert d["impl"] == "reference" and d["metric"] == base["metric"] and d["unit"] == "agent-steps/s"
    for key in ("value", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "cpu_baseline", "e2e", "gpu_launches"):
        assert key in d, key
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]
    # the arm runs exactly the samples it claims, and they fit in the run the driver clocks around it
    assert d["steps"] == 3 and d["warmup"] == 1
    assert 1000.0 <= d["ms_per_step"] < 2500.0 and d["ms_per_step"] * d["steps"] / 1e3 < wall
    # same `config` object as the GPU arm prints for this workload
    import bench
    assert d["config"] == bench.config_dict("c5", bench.WORKLOADS["c5"][2], 1)[0]


def test_bench_config_and_working_set_model():
    import bench
    cfg, flushed = bench.config_dict("c5", 131072, 1)
    assert not flushed and cfg["obs_shape"] == [131072, 4, 3, 11, 11] and cfg["maps"].endswith("first 37 sorted")
    assert bench.config_dict("c2", 4096, 1)[1] and bench.config_dict("c4", 256, 1)[1]        # L2-sized: flushed
    assert not bench.config_dict("c3", 16384, 1)[1] and not bench.config_dict("super", 8192, 1)[1]
    assert bench.config_dict("c5", 131072, 8)[0]["stats_allreduce"].startswith("NCCL")
    assert bench.obs_shape_of("dec", dict(bench.C2_CFG, dijkstra_input=1), 7, 102, 102) == [7, 4, 4, 11, 11]
    assert bench.obs_shape_of("super", bench.WORKLOADS["super"][3], 5, 100, 100) == [5, 7, 100, 100]
