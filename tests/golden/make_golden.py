"""Record golden traces from the live, unmodified reference env (build container only).

Each trace is a session of one reference env object driven with seeded base-4 integer actions
(the only action form both reference envs accept at HEAD — SURVEY.md §3.5).  Map index and start
poses are read back from the reference after every reset and stored, so the replaying side injects
them instead of reproducing numpy's Mersenne-Twister rejection sampling.

Outputs tests/golden/<name>.npz (committed; small).  Usage:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness as rh                                      # noqa: E402
from oracle.trace import OP_RESET, OP_STEP, Recorder, crc_f32, Trace, replay_on_oracle  # noqa: E402
from pin_oracle import dec_cfg, super_cfg, small_maps          # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def snap_reference(env, kind, obs_or_state, reward, done):
    s = dict(pos=np.stack([env._xinds, env._yinds], axis=1).astype(np.int16),
             reward=float(reward), done=int(bool(done)), pct=float(env.percent_covered()),
             done_thresh=float(env._done_thresh), currstep=int(env._currstep))
    if kind == "dec":
        p, W, L = env._pad, env._gridwidth, env._gridlen
        s["obs"] = np.asarray(obs_or_state, dtype=np.float32)
        s["adj"] = env._adjacency_matrix.astype(np.uint8)
        s["ckpt_a"] = env._free_pad[:, p:p + W, p:p + L].astype(np.uint8)
        s["ckpt_b"] = env._obst_pad[:, p:p + W, p:p + L].astype(np.uint8)
        s["ckpt_visited"] = env._visited.astype(np.uint8)
    else:
        st = np.asarray(obs_or_state[0], dtype=np.float32)
        s["state_crc"] = crc_f32(st)
        s["ckpt_a"] = (env._free < 1).astype(np.uint8)
        s["ckpt_b"] = env._observed_obstacles.astype(np.uint8)
        s["ckpt_state"] = st
    return s


def map_index_of(env, kind, maps):
    g = env._grid[1:-1, 1:-1] if kind == "dec" else env._grid
    for i, m in enumerate(maps):
        if m.shape == g.shape and np.array_equal(m, g):
            return i
    raise RuntimeError("reference picked a map that is not in the list")


def record(ns, name, kind, maps, cfg, steps, seed, use_graph=False, episodes=2, ckpt_every=25,
           sentinel_at=None):
    nact = 4 ** cfg["numrobot"]
    np.random.seed(seed)
    with rh.quiet():
        env = (ns.DecGridRL(maps, cfg, use_graph=use_graph) if kind == "dec"
               else ns.SuperGridRL(maps, cfg))
    rec = Recorder(kind, cfg, maps, use_graph=use_graph, ckpt_every=ckpt_every)
    arng = np.random.RandomState(seed + 104729)
    n = cfg["numrobot"]
    zero_start = np.zeros((n, 2), dtype=np.int16)
    for ep in range(episodes):
        np.random.seed(seed * 131 + ep)
        with rh.quiet():
            out = env.reset(False, None)
        mi = map_index_of(env, kind, maps)
        start = np.stack([env._xinds, env._yinds], axis=1).astype(np.int16)
        rec.add(OP_RESET, 0, mi, start, snap_reference(env, kind, out[0], 0.0, False),
                force_ckpt=True)
        for t in range(steps):
            a = int(arng.randint(nact))
            if sentinel_at is not None and ep == 0 and t == sentinel_at:
                a = -1
            with rh.quiet():
                o, r, d = env.step(a)
            if kind == "dec" and use_graph and cfg["allow_comm"]:
                o = o[0]
            last = (t == steps - 1) or (d and a != -1)
            rec.add(OP_STEP, a, mi, zero_start, snap_reference(env, kind, o, r, d),
                    force_ckpt=last)
            if d and a != -1:
                break
    tr = rec.finish()
    path = os.path.join(OUT, name + ".npz")
    tr.save(path)
    # the file must round-trip and the oracle must agree with it
    back = Trace.load(path)
    tol = 1e-6 if cfg.get("mini_map_rad", 0) else 0.0
    nops = replay_on_oracle(back, obs_tol=tol)
    print("%-34s ops=%4d  dones=%2d  %6.1f kB" % (name, nops, int(back.done.sum()),
                                                os.path.getsize(path) / 1e3))


def main():
    ns = rh.load_reference()
    m100 = rh.load_png_dir("bg2_100x100", 4)
    m256 = rh.load_png_dir("bg2_256x256", 1)
    m1073 = rh.load_png_dir("bg2_1073x1073", 1)
    m15 = small_maps()
    circ = rh.load_png_dir("circulartest", 1)
    L = dict(sensor_type="lidar")
    R = record
    # ---- DecGridRL ---------------------------------------------------------------------
    R(ns, "dec_c2_square_100", "dec", m100, dec_cfg(), 300, 101, sentinel_at=150)
    R(ns, "dec_square_15_n3_done", "dec", m15, dec_cfg(numrobot=3, egoradius=3), 200, 102,
      episodes=3)
    R(ns, "dec_square_thresh_incr", "dec", m15,
      dec_cfg(numrobot=2, done_thresh=0.3, done_incr=0.1, maxsteps=60), 120, 103, episodes=5)
    R(ns, "dec_square_n1", "dec", m15, dec_cfg(numrobot=1, egoradius=2), 150, 104)
    R(ns, "dec_square_n6_r3_e7", "dec", m100,
      dec_cfg(numrobot=6, egoradius=7, sensor_config={"range": 3}), 150, 105)
    R(ns, "dec_square_float_penalty", "dec", m15,
      dec_cfg(numrobot=2, collision_penalty=0.3, terminal_reward=2.5, egoradius=4), 150, 106)
    R(ns, "dec_lidar_21x10_100", "dec", m100,
      dec_cfg(numrobot=3, sensor_config={"num_lasers": 21, "range": 10}, **L), 250, 107)
    R(ns, "dec_c3_lidar_256_n8", "dec", m256,
      dec_cfg(numrobot=8, sensor_config={"num_lasers": 21, "range": 10}, **L), 80, 108,
      episodes=1, ckpt_every=40)
    R(ns, "dec_lidar_circular", "dec", circ,
      dec_cfg(numrobot=2, egoradius=4, sensor_config={"num_lasers": 15, "range": 5}, **L),
      150, 109, episodes=3)
    R(ns, "dec_lidar_31x8_15", "dec", m15,
      dec_cfg(numrobot=1, egoradius=5, sensor_config={"num_lasers": 31, "range": 8}, **L),
      150, 110, episodes=3)
    R(ns, "dec_mapsharing_comm12", "dec", m100,
      dec_cfg(numrobot=5, comm_radius=12, allow_comm=1, map_sharing=1), 250, 111,
      use_graph=True)
    R(ns, "dec_single_square_tool", "dec", m100,
      dec_cfg(numrobot=2, single_square_tool=1, egoradius=2), 250, 112)
    R(ns, "dec_dist_reward_15", "dec", m15, dec_cfg(numrobot=3, dist_reward=1, egoradius=3),
      120, 113)
    R(ns, "dec_dijkstra_15", "dec", m15, dec_cfg(numrobot=2, dijkstra_input=1, egoradius=4),
      100, 114)
    R(ns, "dec_minimap_100", "dec", m100, dec_cfg(numrobot=2, mini_map_rad=10), 100, 115)
    R(ns, "dec_c4_square_1073_n16", "dec", m1073, dec_cfg(numrobot=16), 40, 116,
      episodes=1, ckpt_every=1000)
    # ---- SuperGridRL -------------------------------------------------------------------
    R(ns, "super_c1_100", "super", m100, super_cfg(), 120, 201, ckpt_every=30, sentinel_at=60)
    R(ns, "super_int_penalty_episodes", "super", m15,
      super_cfg(free_penalty=0, done_thresh=0.3, done_incr=0.05, numrobot=2), 150, 202,
      episodes=4, ckpt_every=30)
    R(ns, "super_scanning_n5", "super", m100, super_cfg(use_scanning=1, numrobot=5), 80, 203,
      ckpt_every=40)
    R(ns, "super_dist_reward_15", "super", m15, super_cfg(dist_reward=1, numrobot=3), 150, 204,
      ckpt_every=30)
    R(ns, "super_r1_n2_15", "super", m15, super_cfg(senseradius=1, numrobot=2), 200, 205,
      ckpt_every=40)
    R(ns, "super_n9_r3_100", "super", m100, super_cfg(numrobot=9, senseradius=3), 60, 206,
      ckpt_every=30, episodes=1)


if __name__ == "__main__":
    main()
