"""Record end-to-end planner runs from the live reference (build container only): SURVEY.md §7's end-to-end check.

The reference's classical planners (Policies/stc.py:27, bsa.py:64, ba_star.py:122) drive the UNMODIFIED reference
DecGridRL exactly as grid_rl_main.py / Utils/utils.py:generate_episode do, with the env_config of the shipped examples
(Example_Experiments/Non_Learning/{STC,BSA,BA_Star}/Example/config.json).  The planners cannot travel to the GPU box;
their action streams can.  Per testing episode this records the test-map index, the np.random seed the reset ran
under (the drop-in classes consume np.random in the reference's order, so the same seed gives the same start pose),
the start pose, every action the planner chose, and after every step the reward, done flag, percent_covered and the
observation the planner was shown (0/1 layers incl. the dijkstra path layer BSA / BA* read).  tests/ replays the
streams on the oracle (CPU) and on the drop-in DecGridRL (GPU) and asserts the same observations, rewards and the
reference's 100 % coverage (Example/TerminalOutput.txt: "covered 100.0 percent").

Usage:  python tests/golden/make_planner_traces.py      ->  tests/golden/planner_{stc,bsa,ba_star}.npz
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness as rh                                   # noqa: E402

EXAMPLES = {
    # name: (example dir, episodes per test map, number of test maps used, step cap)
    "stc": ("STC", 1, 2, 4000),
    "bsa": ("BSA", 2, 3, 4000),
    "ba_star": ("BA_Star", 2, 3, 4000),
}


def load_example(name):
    d = os.path.join(rh.REFERENCE_ROOT, "Example_Experiments", "Non_Learning", EXAMPLES[name][0], "Example", "config.json")
    with open(d) as f:
        return json.load(f)


def grids_for(grid_config):
    """gridload as grid_rl_main.py:170-176 calls it, but with a SORTED directory listing (gridmaker.py:87 uses raw
    os.listdir order) so the split is reproducible."""
    from Utils.gridmaker import gridload
    if grid_config["grid_dir"] == 0:
        with rh.quiet():
            return gridload(grid_config=None)
    name = os.path.basename(os.path.normpath(grid_config["grid_dir"]))
    maps = rh.load_png_dir(name, grid_config["numgrids"])
    half = len(maps) // 2
    return maps[:half], maps[half:]


def make_policy(name, cfg):
    pc, ec = cfg["policy_config"], cfg["env_config"]
    if name == "stc":
        from Policies.stc import STC
        return STC(pc["internal_grid_rad"])
    if name == "bsa":
        from Policies.bsa import BSA
        return BSA(pc["internal_grid_rad"])
    from Policies.ba_star import BA_Star
    return BA_Star(pc["internal_grid_rad"], ec["egoradius"])


def record(name):
    ns = rh.load_reference()
    cfg = load_example(name)
    env_config = cfg["env_config"]
    train, test = grids_for(cfg["grid_config"])
    _, per_map, n_maps, cap = EXAMPLES[name]
    test = test[:n_maps]
    np.random.seed(12345)
    with rh.quiet():
        env = ns.DecGridRL(train, env_config, use_graph=cfg["policy_config"]["use_graph"], test_set=test)
        policy = make_policy(name, cfg)
    ep_ind, ep_seed, ep_start, ep_len, ep_final = [], [], [], [], []
    actions, rewards, dones, pcts, obs_all = [], [], [], [], []
    for ind in range(len(test)):
        for rep in range(per_map):
            seed = 7000 + 100 * ind + rep
            np.random.seed(seed)
            with rh.quiet():
                state, grid = env.reset(True, ind)
                policy.reset(True, grid)
            ep_ind.append(ind)
            ep_seed.append(seed)
            ep_start.append(np.stack([env._xinds, env._yinds], axis=1))
            obs_all.append(np.asarray(state))
            n = 0
            done = False
            while not done and n < cap:                  # generate_episode, Utils/utils.py:17-42
                with rh.quiet():
                    action = policy.pi(state)
                    state, reward, done = env.step(action)
                if env._currstep == env._test_maxsteps:
                    done = True
                actions.append(int(action))
                rewards.append(float(reward))
                dones.append(bool(done))
                pcts.append(float(env.percent_covered()))
                obs_all.append(np.asarray(state))
                n += 1
            ep_len.append(n)
            ep_final.append(float(env.percent_covered()))
            print("%s: test map %d rep %d: %d steps, covered %.1f %%" % (name, ind, rep, n, 100 * ep_final[-1]))
    obs = np.stack(obs_all)
    assert set(np.unique(obs)) <= {0.0, 1.0}             # all layers of these configs are 0/1: stored as bytes
    out = dict(meta=np.array(json.dumps({"planner": name, "env_config": env_config,
                                         "use_graph": cfg["policy_config"]["use_graph"],
                                         "example": "Example_Experiments/Non_Learning/%s/Example/config.json" % EXAMPLES[name][0]})),
               train=np.stack(train).astype(np.int8), test=np.stack(test).astype(np.int8),
               ep_ind=np.asarray(ep_ind, np.int32), ep_seed=np.asarray(ep_seed, np.int64),
               ep_start=np.asarray(ep_start, np.int16), ep_len=np.asarray(ep_len, np.int32),
               ep_final=np.asarray(ep_final, np.float64), actions=np.asarray(actions, np.int16),
               rewards=np.asarray(rewards, np.float64), dones=np.asarray(dones, np.uint8),
               pcts=np.asarray(pcts, np.float64), obs=obs.astype(np.uint8))
    np.savez_compressed(os.path.join(HERE, "planner_%s.npz" % name), **out)


if __name__ == "__main__":
    for nm in (sys.argv[1:] or list(EXAMPLES)):
        record(nm)
