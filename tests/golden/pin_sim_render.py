"""Pin the oracle's SuperGrid_Sim lookahead and render() frames against the live reference and record
golden vectors (build container only; /root/reference is read, never copied or modified).

Shims, all outside the reference's files:
  * ``numpy.asscalar`` (removed in numpy 1.23, used at sim_env.py:146-147) is re-added as ``a.item()``;
  * ``SuperGrid_Sim.percent_covered`` is called with one argument at sim_env.py:187 but defined with two
    (sim_env.py:248): a subclass supplies the step's own ``grid`` for the missing one;
  * pygame stubs: ``display.set_mode`` returns an object with ``blit``, ``surfarray.make_surface`` keeps
    the array, ``surfarray.array3d`` returns it as uint8 (pygame truncates float pixel values; every
    value render() produces is an integer <= 255).

Outputs: tests/golden/sim_lookahead.npz, tests/golden/render_frames.npz.  Usage: python tests/golden/pin_sim_render.py
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness as rh                                   # noqa: E402
from oracle import coverage_oracle as oc                   # noqa: E402
from pin_oracle import dec_cfg, super_cfg, small_maps      # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


class _Surf:
    def __init__(self, arr):
        self.arr = arr


class _Display:
    def blit(self, *a, **k):
        return None


def install_render_stubs():
    rh._install_stubs()
    pg = sys.modules["pygame"]
    pg.display.set_mode = lambda *a, **k: _Display()
    pg.surfarray.make_surface = lambda arr: _Surf(np.array(arr))
    pg.surfarray.array3d = lambda s: np.asarray(s.arr).astype(np.uint8)


def load_sim():
    install_render_stubs()
    if not hasattr(np, "asscalar"):
        np.asscalar = lambda a: a.item()
    if rh.REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, rh.REFERENCE_ROOT)
    from Environments.sim_env import SuperGrid_Sim

    class Shimmed(SuperGrid_Sim):
        def step(self, state, action, grid):
            self._shim_grid = grid
            return super().step(state, action, grid)

        def percent_covered(self, state, grid=None):
            return super().percent_covered(state, self._shim_grid if grid is None else grid)

    return Shimmed


def pin_sim():
    Sim = load_sim()
    ns = rh.load_reference()
    m100 = rh.load_png_dir("bg2_100x100", 2)
    m15 = small_maps()
    cases = []
    specs = [
        ("c1_100", m100, super_cfg(numrobot=1), 60, 11),
        ("dist_15", m15, super_cfg(numrobot=1, dist_reward=1, senseradius=1), 80, 12),
        ("thresh_15", m15, super_cfg(numrobot=1, done_thresh=0.2, free_penalty=0, senseradius=3), 80, 13),
        ("float_pen_15", m15, super_cfg(numrobot=1, collision_penalty=0.3, free_penalty=0.7,
                                        terminal_reward=2.5), 80, 14),
    ]
    total = 0
    for name, maps, cfg, steps, seed in specs:
        np.random.seed(seed)
        with rh.quiet():
            env = ns.SuperGridRL(maps, cfg)
            sim = Sim(env._obs_dim, cfg)
            (state, _), grid = env.reset(False, None)
        rng = np.random.RandomState(seed + 7)
        mi = [i for i, m in enumerate(maps) if m.shape == grid.shape and np.array_equal(m, grid)][0]
        for t in range(steps):
            # look one step ahead with every action from the live state, then advance the real env
            for a in range(4):
                s_in = (np.array(state, dtype=np.float64), 0)
                want_state, want_r = oc.sim_lookahead(s_in[0], a, grid, cfg)
                with rh.quiet():
                    got_state, got_r = sim.step(s_in, a, grid)
                got_r = float(np.asarray(got_r).reshape(-1)[0])
                assert np.array_equal(got_state, want_state), (name, t, a, "state")
                assert got_r == want_r, (name, t, a, got_r, want_r)
                # the reference mutates the caller's free / observed planes in place
                assert np.array_equal(s_in[0][1], want_state[1]) and np.array_equal(s_in[0][2], want_state[2])
                total += 1
                if t % 4 == 0:
                    cases.append((name, mi, state.astype(np.float32), a, want_r,
                                  zlib.crc32(want_state.astype(np.float32).tobytes()),
                                  want_state.astype(np.float32) if (t % 20 == 0) else None))
            with rh.quiet():
                (state, _), _, d = env.step(int(rng.randint(4)))
            if d:
                break
    print("sim lookahead: %d reference calls match the oracle bit for bit" % total)
    return specs, cases


def save_sim(specs, cases):
    blob = {}
    import json
    meta = []
    for name, maps, cfg, _, _ in specs:
        blob["maps_" + name] = np.stack(maps).astype(np.int8)
        meta.append({"name": name, "cfg": cfg})
    blob["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    for name, _, _, _, _ in specs:
        cs = [c for c in cases if c[0] == name]
        # states are stored as the three bit planes + position (the distance plane is float32)
        blob["st_" + name] = np.stack([c[2] for c in cs])
        blob["mi_" + name] = np.array([c[1] for c in cs], dtype=np.int32)
        blob["act_" + name] = np.array([c[3] for c in cs], dtype=np.uint8)
        blob["rew_" + name] = np.array([c[4] for c in cs], dtype=np.float64)
        blob["crc_" + name] = np.array([c[5] for c in cs], dtype=np.uint32)
        full = [(i, c[6]) for i, c in enumerate(cs) if c[6] is not None]
        blob["fulli_" + name] = np.array([i for i, _ in full], dtype=np.int32)
        blob["full_" + name] = np.stack([f for _, f in full])
    path = os.path.join(OUT, "sim_lookahead.npz")
    np.savez_compressed(path, **blob)
    print("wrote %s (%.1f kB, %d cases)" % (path, os.path.getsize(path) / 1e3, len(cases)))


def pin_render():
    install_render_stubs()
    ns = rh.load_reference()
    m100 = rh.load_png_dir("bg2_100x100", 1)
    m15 = small_maps()
    rect = [np.where(np.random.RandomState(5).rand(20, 37) < 0.15, -1.0, 1.0)]
    frames = {}
    import json
    meta = []
    specs = [("super_100", "super", m100, super_cfg(numrobot=3), 40, 21),
             ("super_scan_15", "super", m15, super_cfg(numrobot=4, use_scanning=1), 30, 22),
             ("super_rect", "super", rect, super_cfg(numrobot=2), 25, 23),
             ("dec_100", "dec", m100, dec_cfg(numrobot=4), 40, 24),
             ("dec_lidar_15", "dec", m15, dec_cfg(numrobot=2, sensor_type="lidar",
                                                  sensor_config={"num_lasers": 15, "range": 5}), 30, 25),
             ("dec_rect", "dec", rect, dec_cfg(numrobot=3), 25, 26)]
    for name, kind, maps, cfg, steps, seed in specs:
        np.random.seed(seed)
        with rh.quiet():
            env = ns.DecGridRL(maps, cfg) if kind == "dec" else ns.SuperGridRL(maps, cfg)
            env.reset(False, None)
        start = np.stack([env._xinds, env._yinds], axis=1).astype(np.int16)
        rng = np.random.RandomState(seed + 3)
        acts = rng.randint(4 ** cfg["numrobot"], size=steps).astype(np.int64)
        for a in acts:
            with rh.quiet():
                _, _, d = env.step(int(a))
            assert not d, "pick a shorter trace"
        with rh.quiet():
            frame = np.asarray(env.render())
        if kind == "dec":
            robots = env._robot_pos_map
            want = oc.render_frame(env._observed_obstacles, env._visited, robots, 1024, True)
        else:
            want = oc.render_frame(env._observed_obstacles, env._free < 1, env.get_pos_image()[0], 1075, False)
        assert frame.shape == want.shape, (name, frame.shape, want.shape)
        assert np.array_equal(np.uint8(frame), want), name
        gq = env._grid[1:-1, 1:-1] if kind == "dec" else env._grid
        mi = [i for i, m in enumerate(maps) if m.shape == gq.shape and np.array_equal(m, gq)][0]
        frames["frame_" + name] = want
        frames["maps_" + name] = np.stack(maps).astype(np.int8)
        frames["start_" + name] = start
        frames["acts_" + name] = acts
        meta.append({"name": name, "kind": kind, "cfg": cfg, "map_index": mi})
        print("render %-14s frame %s matches the oracle" % (name, want.shape))
    frames["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    if CHECK_ONLY:
        return
    path = os.path.join(OUT, "render_frames.npz")
    np.savez_compressed(path, **frames)
    print("wrote %s (%.1f kB)" % (path, os.path.getsize(path) / 1e3))


CHECK_ONLY = "--check" in sys.argv[1:]        # compare only, leave the committed fixtures alone

if __name__ == "__main__":
    specs, cases = pin_sim()
    if not CHECK_ONLY:
        save_sim(specs, cases)
    pin_render()
