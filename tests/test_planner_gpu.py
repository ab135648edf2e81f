"""End-to-end on the GPU: the drop-in DecGridRL (E = 1 view of the batched CUDA env) replays the action streams the
reference's STC / BSA / BA* planners produced on the live reference env (tests/golden/planner_*.npz) — same start
poses from the same np.random seeds, same observations incl. the dijkstra layer, same rewards, and the reference's
100 % coverage for BSA / BA* (Example/TerminalOutput.txt)."""
import pytest

from planner_replay import PLANNERS, replay

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", PLANNERS)
def test_dropin_env_replays_planner_runs(name):
    from marl_coverage_b200 import DecGridRL
    finals = replay(name, lambda train, cfg, ug, test: DecGridRL(train, cfg, use_graph=ug, test_set=test))
    if name in ("bsa", "ba_star"):
        assert finals and all(f == 1.0 for f in finals)
