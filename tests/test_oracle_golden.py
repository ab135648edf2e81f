"""The oracle must reproduce every golden trace recorded from the live reference.

This is what pins oracle/coverage_oracle.py (the reference ships no tests of its own for the env
path — SURVEY.md §4).  Bit-exact everywhere except the bilinear mini-map layers (OpenCV
coefficient precision, tolerance 1e-6 — SURVEY.md row D12)."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, golden_names
from oracle.trace import Trace, replay_on_oracle


@pytest.mark.parametrize("name", golden_names())
def test_oracle_replays_golden(name):
    tr = Trace.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    tol = 1e-6 if tr.config.get("mini_map_rad", 0) else 0.0
    assert replay_on_oracle(tr, obs_tol=tol) == tr.num_ops


def test_pin_report_all_match():
    with open(os.path.join(GOLDEN_DIR, "PIN_REPORT.json")) as f:
        rep = json.load(f)
    assert rep["all_match"] is True
    assert len(rep["cases"]) >= 20
    assert all(c["status"] == "match" for c in rep["cases"].values())


def test_oracle_matches_live_reference():
    """Where the reference is present (the build container, never the GPU box) the pin is re-run live: the
    unmodified reference and the oracle side by side over all pinned configurations, in a subprocess so the
    pygame / matplotlib stubs the reference needs do not leak into this process."""
    import subprocess
    import sys
    ref = os.environ.get("MCOV_REFERENCE_ROOT", "/root/reference")
    if not os.path.isdir(os.path.join(ref, "Environments")):
        pytest.skip("reference not present")
    res = subprocess.run([sys.executable, os.path.join(GOLDEN_DIR, "pin_oracle.py"), "--check"],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert res.stdout.count(" match ") >= 20 and "MISMATCH" not in res.stdout


def test_golden_covers_edge_cases():
    """The fixtures must actually contain the situations Appendix A of SURVEY.md lists."""
    tr = Trace.load(os.path.join(GOLDEN_DIR, "dec_c2_square_100.npz"))
    assert (tr.op_action == -1).any()                      # sentinel step
    assert (tr.reward < 0).any()                           # collisions
    assert tr.obs[0, :, 0].sum() == 0                      # robot layer empty right after reset
    tr = Trace.load(os.path.join(GOLDEN_DIR, "dec_square_thresh_incr.npz"))
    assert tr.done.sum() >= 3 and len(np.unique(tr.done_thresh)) >= 3   # curriculum growth
    tr = Trace.load(os.path.join(GOLDEN_DIR, "dec_square_15_n3_done.npz"))
    assert (tr.pct > 1.0).any()                            # percent_covered sums over robots
    tr = Trace.load(os.path.join(GOLDEN_DIR, "super_c1_100.npz"))
    first_step = int(np.nonzero(tr.op_kind == 1)[0][0])
    assert tr.reward[first_step] <= -4 + 100               # first-step motion penalty present
