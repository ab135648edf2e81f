"""End-to-end pin: the reference's STC / BSA / BA* planners were run on the live reference env with the
shipped example configs and recorded (tests/golden/make_planner_traces.py); here the oracle replays their action
streams and must show the planners the same observations — incl. the dijkstra path layer BSA / BA* steer by — and reach
the reference's 100 % coverage (Example_Experiments/Non_Learning/BA_Star/Example/TerminalOutput.txt:172-173)."""
import pytest

from oracle import coverage_oracle as oc
from planner_replay import PLANNERS, load, replay


@pytest.mark.parametrize("name", PLANNERS)
def test_oracle_replays_planner_runs(name):
    finals = replay(name, lambda train, cfg, ug, test: oc.OracleDecGridRL(train, cfg, use_graph=ug, test_set=test))
    if name in ("bsa", "ba_star"):
        assert finals and all(f == 1.0 for f in finals)


def test_traces_hold_what_the_planners_need():
    meta, z = load("ba_star")
    assert meta["env_config"]["dijkstra_input"] == 1 and z["obs"].shape[2] == 4      # layer 3 = dijkstra path
    assert z["obs"][:, :, 3].any()
    meta, z = load("stc")
    assert (z["actions"] == -1).any()                                             # STC ends with the sentinel
