"""Replay of the planner runs recorded from the live reference (tests/golden/make_planner_traces.py) on any env
object with the reference's surface: the oracle on CPU, the drop-in DecGridRL on the GPU."""
import contextlib
import io
import json
import os

import numpy as np

from conftest import GOLDEN_DIR

PLANNERS = ("stc", "bsa", "ba_star")


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, "planner_%s.npz" % name))
    meta = json.loads(str(z["meta"]))
    return meta, z


def replay(name, make_env):
    """make_env(train_set, env_config, use_graph, test_set) -> env.  Walks every recorded testing episode the way
    Utils/utils.py:generate_episode does, feeding the planner's recorded actions; asserts start pose, every
    observation the planner was shown, reward, done and percent_covered.  Returns the final coverage per episode."""
    meta, z = load(name)
    cfg = meta["env_config"]
    train = [m.astype(np.float64) for m in z["train"]]
    test = [m.astype(np.float64) for m in z["test"]]
    with contextlib.redirect_stdout(io.StringIO()):
        np.random.seed(12345)
        env = make_env(train, cfg, bool(meta["use_graph"]), test)
    k = 0                                                   # running index into the per-step arrays
    o = 0                                                   # running index into obs (one extra row per episode)
    finals = []
    for ep in range(len(z["ep_ind"])):
        np.random.seed(int(z["ep_seed"][ep]))
        with contextlib.redirect_stdout(io.StringIO()):
            state, grid = env.reset(True, int(z["ep_ind"][ep]))
        start = np.stack([env._xinds, env._yinds], axis=1)
        assert np.array_equal(start, z["ep_start"][ep]), "%s episode %d: start pose %s vs %s" % (
            name, ep, start.tolist(), z["ep_start"][ep].tolist())
        assert np.array_equal(np.pad(test[int(z["ep_ind"][ep])], 1, constant_values=-1), grid)
        assert np.array_equal(np.asarray(state), z["obs"][o].astype(np.float64)), "%s episode %d: reset obs" % (name, ep)
        o += 1
        for t in range(int(z["ep_len"][ep])):
            with contextlib.redirect_stdout(io.StringIO()):
                state, reward, done = env.step(int(z["actions"][k]))
            if env._currstep == env._test_maxsteps:
                done = True
            tag = "%s episode %d step %d (action %d)" % (name, ep, t, int(z["actions"][k]))
            assert np.array_equal(np.asarray(state), z["obs"][o].astype(np.float64)), "obs, " + tag
            assert float(reward) == float(z["rewards"][k]), "reward %r vs %r, %s" % (reward, z["rewards"][k], tag)
            assert bool(done) == bool(z["dones"][k]), "done, " + tag
            assert float(env.percent_covered()) == float(z["pcts"][k]), "percent_covered, " + tag
            k += 1
            o += 1
        assert bool(z["dones"][k - 1])                       # every recorded episode ran to its end
        finals.append(float(env.percent_covered()))
        assert finals[-1] == float(z["ep_final"][ep])
    assert k == len(z["actions"]) and o == len(z["obs"])
    return finals
