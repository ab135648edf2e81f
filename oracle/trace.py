"""Golden-trace container for env-step parity (TEST INFRASTRUCTURE).

A trace is one *session* of one env object: a list of ops (``reset`` with an injected map index
and start poses, or ``step`` with a base-4 joint action / the -1 sentinel) together with what the
reference returned / held after each op.  ``tests/golden/make_golden.py`` records them from the live
reference; ``tests/`` replays them against the oracle (CPU) and the CUDA env (GPU).

File format: ``np.savez_compressed`` with the keys documented in ``Trace``.
"""
import json
import zlib

import numpy as np

OP_RESET, OP_STEP = 0, 1


def crc_f32(a):
    """CRC32 of the float32 little-endian bytes of ``a`` — a bit-exactness fingerprint for
    tensors too large to store per step (SuperGridRL states)."""
    return zlib.crc32(np.ascontiguousarray(a, dtype="<f4").tobytes()) & 0xFFFFFFFF


class Trace:
    """
    kind          'dec' | 'super'
    config        env_config dict (reference keys)
    use_graph     bool (Dec only)
    maps          int8 [M, W, L]  (-1 / 0 / +1), UN-bordered, the env's train_set
    op_kind       uint8 [K]       OP_RESET / OP_STEP
    op_action     int64 [K]       joint base-4 action (step) — -1 is the sentinel
    op_map        int32 [K]       map index in force after the op
    op_start      int16 [K, N, 2] start poses (reset ops; env coordinates, i.e. bordered for Dec)
    pos           int16 [K, N, 2] positions after the op
    reward        float64 [K]     (0 for resets)
    done          uint8 [K]
    pct           float64 [K]     percent_covered() after the op
    done_thresh   float64 [K]
    currstep      int32 [K]
    obs           Dec: float32 [K, N, C, h, w];  Super: absent
    state_crc     Super: uint32 [K] CRC32 of the float32 state tensor
    adj           Dec: uint8 [K, N, N]
    ckpt_ops      int32 [Q]       op indices with full planes stored
    ckpt_a        Dec: uint8 [Q, N, W', L'] per-robot free planes;  Super: uint8 [Q, W, L] covered
    ckpt_b        Dec: uint8 [Q, N, W', L'] per-robot obstacle planes; Super: [Q, W, L] observed obst.
    ckpt_visited  Dec: uint8 [Q, W', L']
    ckpt_state    Super: float32 [Q, C, W, L] full state tensors
    """

    FIELDS = ("maps", "op_kind", "op_action", "op_map", "op_start", "pos", "reward", "done", "pct",
              "done_thresh", "currstep", "obs", "state_crc", "adj", "ckpt_ops", "ckpt_a",
              "ckpt_b", "ckpt_visited", "ckpt_state")

    def __init__(self, kind, config, use_graph=False):
        self.kind = kind
        self.config = config
        self.use_graph = bool(use_graph)
        for f in self.FIELDS:
            setattr(self, f, None)

    def save(self, path):
        d = {"meta": np.array(json.dumps({"kind": self.kind, "config": self.config,
                                          "use_graph": self.use_graph}))}
        for f in self.FIELDS:
            v = getattr(self, f)
            if v is not None:
                d[f] = v
        np.savez_compressed(path, **d)

    @staticmethod
    def load(path):
        z = np.load(path, allow_pickle=False)
        meta = json.loads(str(z["meta"]))
        t = Trace(meta["kind"], meta["config"], meta["use_graph"])
        for f in Trace.FIELDS:
            if f in z.files:
                setattr(t, f, z[f])
        return t

    @property
    def num_ops(self):
        return int(self.op_kind.shape[0])

    def float_maps(self):
        """The train_set list (float64 grids) the reference env was built with."""
        return [m.astype(np.float64) for m in self.maps]


class Recorder:
    """Accumulates per-op rows; adapters supply the env-specific getters."""

    def __init__(self, kind, config, maps, use_graph=False, ckpt_every=25):
        self.t = Trace(kind, config, use_graph)
        self.t.maps = np.stack([np.asarray(m) for m in maps]).astype(np.int8)
        self.rows = {k: [] for k in ("op_kind", "op_action", "op_map", "op_start", "pos", "reward",
                                     "done", "pct", "done_thresh", "currstep", "obs", "state_crc",
                                     "adj")}
        self.ck = {k: [] for k in ("ckpt_ops", "ckpt_a", "ckpt_b", "ckpt_visited", "ckpt_state")}
        self.ckpt_every = ckpt_every
        self.n = 0

    def add(self, op_kind, action, map_idx, start, snap, force_ckpt=False):
        r = self.rows
        r["op_kind"].append(op_kind)
        r["op_action"].append(action)
        r["op_map"].append(map_idx)
        r["op_start"].append(start)
        for k in ("pos", "reward", "done", "pct", "done_thresh", "currstep"):
            r[k].append(snap[k])
        for k in ("obs", "state_crc", "adj"):
            if k in snap:
                r[k].append(snap[k])
        if force_ckpt or self.n % self.ckpt_every == 0:
            self.ck["ckpt_ops"].append(self.n)
            for k in ("ckpt_a", "ckpt_b", "ckpt_visited", "ckpt_state"):
                if k in snap:
                    self.ck[k].append(snap[k])
        self.n += 1

    def finish(self):
        t, r = self.t, self.rows
        t.op_kind = np.asarray(r["op_kind"], dtype=np.uint8)
        t.op_action = np.asarray(r["op_action"], dtype=np.int64)
        t.op_map = np.asarray(r["op_map"], dtype=np.int32)
        t.op_start = np.asarray(r["op_start"], dtype=np.int16)
        t.pos = np.asarray(r["pos"], dtype=np.int16)
        t.reward = np.asarray(r["reward"], dtype=np.float64)
        t.done = np.asarray(r["done"], dtype=np.uint8)
        t.pct = np.asarray(r["pct"], dtype=np.float64)
        t.done_thresh = np.asarray(r["done_thresh"], dtype=np.float64)
        t.currstep = np.asarray(r["currstep"], dtype=np.int32)
        if r["obs"]:
            t.obs = np.asarray(r["obs"], dtype=np.float32)
        if r["state_crc"]:
            t.state_crc = np.asarray(r["state_crc"], dtype=np.uint32)
        if r["adj"]:
            t.adj = np.asarray(r["adj"], dtype=np.uint8)
        t.ckpt_ops = np.asarray(self.ck["ckpt_ops"], dtype=np.int32)
        for k, dt in (("ckpt_a", np.uint8), ("ckpt_b", np.uint8), ("ckpt_visited", np.uint8),
                      ("ckpt_state", np.float32)):
            if self.ck[k]:
                setattr(t, k, np.asarray(self.ck[k], dtype=dt))
        return t


# ---------------------------------------------------------------------------------------------
# snapshots of the ORACLE env (the reference-side twin lives in tests/golden/make_golden.py)
# ---------------------------------------------------------------------------------------------

def snap_oracle(env, kind, obs_or_state, reward, done):
    s = dict(pos=np.stack([env._xinds, env._yinds], axis=1).astype(np.int16),
             reward=float(reward), done=int(bool(done)), pct=float(env.percent_covered()),
             done_thresh=float(env._done_thresh), currstep=int(env._currstep))
    if kind == "dec":
        s["obs"] = np.asarray(obs_or_state, dtype=np.float32)
        s["adj"] = env._adjacency_matrix.astype(np.uint8)
        s["ckpt_a"] = env._free.astype(np.uint8)
        s["ckpt_b"] = env._obst.astype(np.uint8)
        s["ckpt_visited"] = env._visited.astype(np.uint8)
    else:
        st = np.asarray(obs_or_state[0], dtype=np.float32)
        s["state_crc"] = crc_f32(st)
        s["ckpt_a"] = env._covered.astype(np.uint8)
        s["ckpt_b"] = env._observed_obstacles.astype(np.uint8)
        s["ckpt_state"] = st
    return s


def compare_snap(trace, k, snap, obs_tol=0.0, reward_rtol=0.0, where=""):
    """Raises AssertionError if ``snap`` (from any implementation) differs from op ``k`` of the
    golden trace.  Integer/bit fields exact; ``obs`` exact unless ``obs_tol``; reward exact unless
    ``reward_rtol``."""
    tag = "%s op %d" % (where, k)
    assert np.array_equal(trace.pos[k], snap["pos"]), "%s: positions %s vs %s" % (
        tag, trace.pos[k].tolist(), np.asarray(snap["pos"]).tolist())
    assert int(trace.done[k]) == int(snap["done"]), "%s: done" % tag
    if reward_rtol == 0.0:
        assert trace.reward[k] == snap["reward"], "%s: reward %r vs %r" % (
            tag, trace.reward[k], snap["reward"])
    else:
        assert abs(trace.reward[k] - snap["reward"]) <= reward_rtol * max(1.0, abs(trace.reward[k])), \
            "%s: reward %r vs %r" % (tag, trace.reward[k], snap["reward"])
    assert trace.pct[k] == snap["pct"], "%s: percent_covered" % tag
    assert trace.done_thresh[k] == snap["done_thresh"], "%s: done_thresh" % tag
    assert int(trace.currstep[k]) == int(snap["currstep"]), "%s: currstep" % tag
    if trace.obs is not None and "obs" in snap:
        a, b = trace.obs[k], np.asarray(snap["obs"], dtype=np.float32)
        assert a.shape == b.shape, "%s: obs shape %s vs %s" % (tag, a.shape, b.shape)
        if obs_tol == 0.0:
            assert np.array_equal(a, b), "%s: obs differ at %s" % (
                tag, np.argwhere(a != b)[:5].tolist())
        else:
            assert np.abs(a - b).max() <= obs_tol, "%s: obs err %g" % (tag, np.abs(a - b).max())
    if trace.state_crc is not None and "state_crc" in snap:
        assert int(trace.state_crc[k]) == int(snap["state_crc"]), "%s: state CRC" % tag
    if trace.adj is not None and "adj" in snap:
        assert np.array_equal(trace.adj[k], snap["adj"]), "%s: adjacency" % tag
    q = np.nonzero(trace.ckpt_ops == k)[0]
    if q.size:
        q = int(q[0])
        for name in ("ckpt_a", "ckpt_b", "ckpt_visited", "ckpt_state"):
            ref = getattr(trace, name)
            if ref is not None and name in snap:
                got = np.asarray(snap[name])
                if name == "ckpt_state" and obs_tol:
                    assert np.abs(ref[q] - got).max() <= obs_tol, "%s: %s" % (tag, name)
                else:
                    assert np.array_equal(ref[q], got), "%s: %s differs" % (tag, name)


def replay_on_oracle(trace, obs_tol=0.0):
    """Replays a golden trace through oracle/coverage_oracle.py; asserts equality per op."""
    from . import coverage_oracle as oc
    maps = trace.float_maps()
    # the ctor performs a throw-away random reset; seed it so a failure is reproducible
    state = np.random.get_state()
    np.random.seed(0)
    try:
        if trace.kind == "dec":
            env = oc.OracleDecGridRL(maps, trace.config, use_graph=trace.use_graph)
        else:
            env = oc.OracleSuperGridRL(maps, trace.config)
    finally:
        np.random.set_state(state)
    for k in range(trace.num_ops):
        if trace.op_kind[k] == OP_RESET:
            out = env.reset(False, None, start_xy=trace.op_start[k], map_index=int(trace.op_map[k]))
            snap = snap_oracle(env, trace.kind, out[0], 0.0, False)
        else:
            o, r, d = env.step(int(trace.op_action[k]))
            if trace.kind == "dec" and trace.use_graph and trace.config["allow_comm"]:
                o = o[0]
            snap = snap_oracle(env, trace.kind, o, r, d)
        compare_snap(trace, k, snap, obs_tol=obs_tol, where="oracle")
    return trace.num_ops
