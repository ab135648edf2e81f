"""Pin oracle/coverage_oracle.py against the live reference (build container only).

Runs the unmodified reference env and the oracle side by side from the same ``np.random`` seed and
the same integer (base-4) action stream, and compares after every reset/step:
positions, reward, done flag, observation/state tensors, the per-robot free/obstacle planes,
visited / observed-obstacle planes, adjacency, percent_covered and done_thresh.

Writes tests/golden/PIN_REPORT.json (committed).  Usage:  python tests/golden/pin_oracle.py
``--check`` runs the same comparison without touching the report (tests/test_oracle_golden.py uses it wherever
/root/reference is present).
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness as rh                                   # noqa: E402
from oracle import coverage_oracle as oc                   # noqa: E402


def dec_cfg(**kw):
    c = dict(numrobot=4, maxsteps=1000, collision_penalty=5, done_thresh=1, done_incr=0,
             terminal_reward=30, dist_reward=0, train_maxsteps=1000, test_maxsteps=1000,
             egoradius=5, mini_map_rad=0, comm_radius=0, allow_comm=0, map_sharing=0,
             single_square_tool=0, dijkstra_input=0, sensor_type="square_sensor",
             sensor_config={"range": 2})
    c.update(kw)
    return c


def super_cfg(**kw):
    c = dict(numrobot=4, train_maxsteps=1000, test_maxsteps=1000, collision_penalty=5,
             senseradius=2, free_penalty=0.2, done_thresh=1, done_incr=0, terminal_reward=30,
             dist_reward=0, use_scanning=0)
    c.update(kw)
    return c


def small_maps():
    """The six 15x15 training maps the reference builds when grid_config is None
    (Utils/gridmaker.py:46-80) plus the 12x12 circular test map, regenerated through the
    reference's own loader so nothing is transcribed by hand."""
    sys.path.insert(0, rh.REFERENCE_ROOT)
    rh._install_stubs()
    from Utils.gridmaker import gridload
    with rh.quiet():
        train, test = gridload(None)
    return train + test


def valued(maps, seed, worths=(0.0, 0.25, 1.0, 2.5, 3.0, 0.1)):
    """The same maps with free cells worth 0 / 0.25 / 1 / 2.5 / 3 / 0.1 instead of 1: the reference pays
    ``reward += self._grid[j][k]`` (super_grid_rl.py:189), counts ``grid > 0`` and walks on ``grid >= 0``."""
    rng = np.random.RandomState(seed)
    out = []
    for m in maps:
        m = np.asarray(m, dtype=np.float64)
        worth = rng.choice(list(worths), size=m.shape)
        out.append(np.where(m > 0, worth, m * rng.choice([1.0, 2.0], size=m.shape)))
    return out


class Mismatch(Exception):
    pass


def _eq(name, a, b, tol=None):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.shape != b.shape:
        raise Mismatch("%s shape %s vs %s" % (name, a.shape, b.shape))
    if tol is None:
        if not np.array_equal(a, b):
            raise Mismatch("%s differs (max abs %g)" % (name, np.abs(a - b).max()))
        return 0.0
    err = float(np.abs(a - b).max()) if a.size else 0.0
    if err > tol:
        raise Mismatch("%s differs by %g > %g" % (name, err, tol))
    return err


def compare_dec(ref, orc, obs_r, obs_o, tol_obs):
    p = ref._pad
    W, L = ref._gridwidth, ref._gridlen
    _eq("xinds", ref._xinds, orc._xinds)
    _eq("yinds", ref._yinds, orc._yinds)
    _eq("free", ref._free_pad[:, p:p + W, p:p + L], orc._free)
    _eq("obst", ref._obst_pad[:, p:p + W, p:p + L], orc._obst)
    # nothing may ever land in the padding ring
    if ref._free_pad.sum() != ref._free_pad[:, p:p + W, p:p + L].sum():
        raise Mismatch("reference wrote free bits into the padding")
    _eq("visited", ref._visited, orc._visited)
    _eq("observed_obstacles", ref._observed_obstacles, orc._observed_obstacles)
    _eq("adjacency", ref._adjacency_matrix, orc._adjacency_matrix)
    _eq("percent_covered", ref.percent_covered(), orc.percent_covered())
    _eq("done_thresh", ref._done_thresh, orc._done_thresh)
    _eq("currstep", ref._currstep, orc._currstep)
    return _eq("obs", obs_r, obs_o, tol_obs)


def compare_super(ref, orc, st_r, st_o):
    _eq("xinds", ref._xinds, orc._xinds)
    _eq("yinds", ref._yinds, orc._yinds)
    _eq("free", ref._free, (~orc._covered).astype(float))
    _eq("observed_obstacles", ref._observed_obstacles, orc._observed_obstacles)
    _eq("percent_covered", ref.percent_covered(), orc.percent_covered())
    _eq("done_thresh", ref._done_thresh, orc._done_thresh)
    _eq("state", st_r[0], st_o[0])
    _eq("currstep", st_r[1], st_o[1])


def run_case(ns, kind, maps, cfg, steps, seed, tol_obs=None, use_graph=False, episodes=2):
    nact = 4 ** cfg["numrobot"]
    np.random.seed(seed)
    with rh.quiet():
        ref = (ns.DecGridRL(maps, cfg, use_graph=use_graph) if kind == "dec"
               else ns.SuperGridRL(maps, cfg))
    np.random.seed(seed)
    orc = (oc.OracleDecGridRL(maps, cfg, use_graph=use_graph) if kind == "dec"
           else oc.OracleSuperGridRL(maps, cfg))
    arng = np.random.RandomState(seed + 7919)
    worst_r, worst_o, nsteps, ndone = 0.0, 0.0, 0, 0
    for ep in range(episodes):
        np.random.seed(seed * 31 + ep)
        with rh.quiet():
            out_r = ref.reset(False, None)
        np.random.seed(seed * 31 + ep)
        out_o = orc.reset(False, None)
        _eq("grid", out_r[1], out_o[1])
        if kind == "dec":
            worst_o = max(worst_o, compare_dec(ref, orc, out_r[0], out_o[0], tol_obs))
        else:
            compare_super(ref, orc, out_r[0], out_o[0])
        for t in range(steps):
            a = int(arng.randint(nact))
            if t == steps // 2 and ep == 0:
                a = -1                                   # sentinel path
            with rh.quiet():
                s_r, r_r, d_r = ref.step(a)
            s_o, r_o, d_o = orc.step(a)
            if bool(d_r) != bool(d_o):
                raise Mismatch("done differs at step %d" % t)
            if r_r != r_o:
                rel = abs(r_r - r_o) / max(abs(r_r), 1e-300)
                worst_r = max(worst_r, rel)
                if rel > 1e-12:
                    raise Mismatch("reward %r vs %r at step %d" % (r_r, r_o, t))
            if kind == "dec":
                if use_graph and cfg["allow_comm"]:
                    _eq("adj_out", s_r[1], s_o[1])
                    s_r, s_o = s_r[0], s_o[0]
                worst_o = max(worst_o, compare_dec(ref, orc, s_r, s_o, tol_obs))
            else:
                compare_super(ref, orc, s_r, s_o)
            nsteps += 1
            if d_r and a != -1:
                ndone += 1
                break
    return dict(steps=nsteps, episodes_done=ndone, reward_max_rel=worst_r, obs_max_abs=worst_o)


def main():
    ns = rh.load_reference()
    m100 = rh.load_png_dir("bg2_100x100", 6)
    m256 = rh.load_png_dir("bg2_256x256", 2)
    m15 = small_maps()
    circ = rh.load_png_dir("circulartest", 1)
    cases = [
        ("dec_square_c2", "dec", m100, dec_cfg(), 400, 1, {}),
        ("dec_square_15x15_n3", "dec", m15, dec_cfg(numrobot=3, egoradius=3), 300, 2, {}),
        ("dec_square_done_lowthresh", "dec", m15,
         dec_cfg(numrobot=2, done_thresh=0.3, done_incr=0.1, maxsteps=60), 200, 3,
         {"episodes": 4}),
        ("dec_lidar_21x10", "dec", m100,
         dec_cfg(numrobot=3, sensor_type="lidar", sensor_config={"num_lasers": 21, "range": 10}),
         300, 4, {}),
        ("dec_lidar_31x8_256", "dec", m256,
         dec_cfg(numrobot=8, sensor_type="lidar", sensor_config={"num_lasers": 31, "range": 8}),
         60, 5, {"episodes": 1}),
        ("dec_lidar_circular", "dec", circ,
         dec_cfg(numrobot=2, egoradius=4,
                 sensor_type="lidar", sensor_config={"num_lasers": 15, "range": 5}), 200, 6, {}),
        ("dec_mapsharing_comm12", "dec", m100,
         dec_cfg(numrobot=5, comm_radius=12, allow_comm=1, map_sharing=1), 300, 7,
         {"use_graph": True}),
        ("dec_single_square_tool", "dec", m100,
         dec_cfg(numrobot=2, single_square_tool=1, egoradius=2), 300, 8, {}),
        ("dec_dist_reward", "dec", m15, dec_cfg(numrobot=3, dist_reward=1, egoradius=3),
         150, 9, {}),
        ("dec_dist_reward_100", "dec", m100, dec_cfg(numrobot=2, dist_reward=1), 60, 10,
         {"episodes": 1}),
        ("dec_dijkstra", "dec", m15,
         dec_cfg(numrobot=2, dijkstra_input=1, egoradius=4), 120, 11, {}),
        ("dec_minimap", "dec", m100, dec_cfg(numrobot=2, mini_map_rad=10), 150, 12,
         {"tol_obs": 1e-6}),
        ("dec_all_layers", "dec", m15,
         dec_cfg(numrobot=2, mini_map_rad=6, egoradius=3, dist_reward=1, dijkstra_input=1),
         80, 13, {"tol_obs": 1e-6}),
        ("dec_n1_float_penalty", "dec", m15,
         dec_cfg(numrobot=1, collision_penalty=0.3, terminal_reward=2.5), 200, 14, {}),
        ("super_c1", "super", m100, super_cfg(), 400, 21, {}),
        ("super_int_penalty", "super", m100,
         super_cfg(free_penalty=0, done_thresh=0.3, done_incr=0.05), 400, 22, {"episodes": 3}),
        ("super_scanning", "super", m100, super_cfg(use_scanning=1, numrobot=5), 300, 23, {}),
        ("super_dist_reward", "super", m15, super_cfg(dist_reward=1, numrobot=3), 200, 24, {}),
        ("super_15x15_r1", "super", m15, super_cfg(senseradius=1, numrobot=2), 300, 25, {}),
        ("super_n9", "super", m100, super_cfg(numrobot=9, senseradius=3), 150, 26, {}),
        ("super_cell_values", "super", valued(m15, 1), super_cfg(numrobot=3, senseradius=1, done_thresh=0.6), 250, 27,
         {"episodes": 3}),
        ("super_cell_values_100_scanning", "super", valued(m100[:2], 2), super_cfg(use_scanning=1, free_penalty=0.37),
         150, 28, {"episodes": 1}),
        # DecGridRL: the square sensor stores clip(cell, 0, 1), so magnitudes >= 1 behave like +-1 (and 0 = walkable,
        # never known-free); the reward stays a count
        ("dec_cell_values_clip", "dec", valued(m15, 3, (0.0, 1.0, 2.5, 3.0)),
         dec_cfg(numrobot=2, egoradius=3, done_thresh=0.5), 150, 29, {}),
        ("dec_cell_values_clip_lidar", "dec", valued(m15, 4, (0.0, 1.0, 2.5, 3.0)),
         dec_cfg(numrobot=2, egoradius=3, sensor_type="lidar", sensor_config={"num_lasers": 9, "range": 4}), 100, 30, {}),
    ]
    report = {"reference_root": rh.REFERENCE_ROOT, "numpy": np.__version__,
              "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), "cases": {}}
    ok = True
    for name, kind, maps, cfg, steps, seed, extra in cases:
        t0 = time.time()
        try:
            res = run_case(ns, kind, maps, cfg, steps, seed, **extra)
            res["status"] = "match"
        except Mismatch as e:
            res = {"status": "MISMATCH", "detail": str(e)}
            ok = False
        res["config"] = cfg
        res["seconds"] = round(time.time() - t0, 2)
        report["cases"][name] = res
        print(name, res["status"], {k: v for k, v in res.items() if k not in ("config",)})
    report["all_match"] = ok
    if "--check" not in sys.argv[1:]:
        with open(os.path.join(ROOT, "tests", "golden", "PIN_REPORT.json"), "w") as f:
            json.dump(report, f, indent=1, sort_keys=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
