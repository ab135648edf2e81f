import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_names(kind=None):
    out = []
    for f in sorted(os.listdir(GOLDEN_DIR)):
        if not f.endswith(".npz") or not f.startswith(("dec_", "super_")):
            continue                                   # env traces only (replay_ring.npz is a different fixture)
        if kind is None or f.startswith(kind + "_"):
            out.append(f[:-4])
    return out


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN_DIR
