"""Golden vectors for the replay-buffer rows, recorded from the reference's own ReplayBuffer
(Policies/replaybuffer.py, pure numpy) in the build container.  Output: tests/golden/replay_ring.npz."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)
import ref_harness as rh                                      # noqa: E402


def main():
    sys.path.insert(0, rh.REFERENCE_ROOT)
    from Policies.replaybuffer import ReplayBuffer
    rng = np.random.default_rng(42)
    obs_dim, act_dim, size, n = (3, 5, 5), 2, 23, 61           # wraps the ring more than twice
    buf = ReplayBuffer(obs_dim, act_dim, size)
    st = rng.integers(0, 2, size=(n,) + obs_dim).astype(np.float32)
    nx = rng.integers(0, 2, size=(n,) + obs_dim).astype(np.float32)
    ac = rng.integers(0, 4, size=(n, act_dim)).astype(np.float32)
    rw = (rng.integers(-5, 31, size=n) + rng.integers(0, 2, size=n) * 0.5).astype(np.float32)
    dn = (rng.random(n) < 0.1).astype(np.float32)
    for k in range(n):
        buf.addtransition(st[k], ac[k], rw[k], nx[k], dn[k])
    idx = rng.integers(0, buf._size, size=17)
    np.random.seed(5)
    got = buf.samplebatch(9)
    np.random.seed(5)
    idx2 = np.random.randint(0, buf._size, size=9)
    assert np.array_equal(got[0], buf._state[idx2])
    out = os.path.join(ROOT, "tests", "golden", "replay_ring.npz")
    np.savez_compressed(out, st=st, nx=nx, ac=ac, rw=rw, dn=dn, size=size, idx=idx, ptr=buf._ptr, n_size=buf._size,
                        ring_state=buf._state, ring_next=buf._nextstate, ring_action=buf._action,
                        ring_reward=buf._reward, ring_done=buf._done, seed5_state=got[0], seed5_reward=got[2])
    print("wrote", out, os.path.getsize(out))


if __name__ == "__main__":
    main()
