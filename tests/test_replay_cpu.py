"""The oracle ring buffer against golden vectors recorded from the reference's ReplayBuffer
(tests/golden/make_golden_replay.py), and — when /root/reference is present — against the reference live."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from oracle.coverage_oracle import OracleReplayBuffer


def _fill(buf, g, lo, hi):
    for k in range(lo, hi):
        buf.addtransition(g["st"][k], g["ac"][k], g["rw"][k], g["nx"][k], g["dn"][k])


def test_oracle_ring_matches_reference_golden():
    g = np.load(os.path.join(GOLDEN_DIR, "replay_ring.npz"))
    buf = OracleReplayBuffer((3, 5, 5), 2, int(g["size"]))
    _fill(buf, g, 0, g["st"].shape[0])
    assert buf._ptr == int(g["ptr"]) and buf._size == int(g["n_size"])
    for a, b in ((buf._state, "ring_state"), (buf._nextstate, "ring_next"), (buf._action, "ring_action"),
                 (buf._reward, "ring_reward"), (buf._done, "ring_done")):
        assert np.array_equal(a, g[b]), b
    np.random.seed(5)
    idx = np.random.randint(0, buf._size, size=9)
    out = buf.gather(idx)
    assert np.array_equal(out[0], g["seed5_state"]) and np.array_equal(out[2], g["seed5_reward"])


def test_oracle_ring_matches_live_reference():
    ref = os.environ.get("MCOV_REFERENCE_ROOT", "/root/reference")
    if not os.path.isdir(os.path.join(ref, "Policies")):
        pytest.skip("reference not present (GPU box)")
    sys.path.insert(0, ref)
    from Policies.replaybuffer import ReplayBuffer
    rng = np.random.default_rng(1)
    a, b = ReplayBuffer((4, 3), 3, 11), OracleReplayBuffer((4, 3), 3, 11)
    for _ in range(40):
        row = (rng.random((4, 3)).astype(np.float32), rng.integers(0, 4, 3), float(rng.integers(-5, 30)),
               rng.random((4, 3)).astype(np.float32), bool(rng.random() < 0.2))
        a.addtransition(*row)
        b.addtransition(*row)
    assert a._ptr == b._ptr and a._size == b._size
    for n in ("_state", "_action", "_reward", "_nextstate", "_done"):
        assert np.array_equal(getattr(a, n), getattr(b, n)), n
