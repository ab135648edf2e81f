"""Import the UNMODIFIED reference env from /root/reference (this container only).

Test infrastructure, not product code.  The reference (ExistentialRobotics/MARL-Coverage)
imports ``pygame`` and ``matplotlib`` at module top level
(Environments/super_grid_rl.py:11,14; Environments/dec_grid_rl.py:10,15) but only uses them
for a window and plots; neither is installed here, so two empty stub modules are placed in
``sys.modules`` before the import.  Nothing under /root/reference is copied or modified.

/root/reference does not exist on the GPU box: only ``tests/golden/make_golden.py`` and
``tests/golden/pin_oracle.py`` (run here, outputs committed under tests/golden/) import this file.
"""
import contextlib
import io
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("MCOV_REFERENCE_ROOT", "/root/reference")


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "Environments"))


def _install_stubs():
    if "pygame" not in sys.modules:
        pg = types.ModuleType("pygame")
        pg.init = lambda *a, **k: None
        disp = types.ModuleType("pygame.display")
        disp.set_mode = lambda *a, **k: None
        disp.update = lambda *a, **k: None
        pg.display = disp
        surf = types.ModuleType("pygame.surfarray")
        pg.surfarray = surf
        sys.modules["pygame"] = pg
        sys.modules["pygame.display"] = disp
        sys.modules["pygame.surfarray"] = surf
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt


def load_reference():
    """Returns a namespace with the reference classes (SuperGridRL, DecGridRL, sensors...)."""
    if not reference_available():
        raise RuntimeError("reference not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    ns = types.SimpleNamespace()
    from Environments.super_grid_rl import SuperGridRL
    from Environments.dec_grid_rl import DecGridRL
    from Environments.Sensors.squaresensor import SquareSensor
    from Environments.Sensors.lidar import LidarSensor
    from Environments.dijkstra import dijkstra_path_map, dijkstra_cost_map
    from Action_Spaces.discrete import Discrete
    ns.SuperGridRL = SuperGridRL
    ns.DecGridRL = DecGridRL
    ns.SquareSensor = SquareSensor
    ns.LidarSensor = LidarSensor
    ns.dijkstra_path_map = dijkstra_path_map
    ns.dijkstra_cost_map = dijkstra_cost_map
    ns.Discrete = Discrete
    return ns


@contextlib.contextmanager
def quiet():
    """The reference prints from done() and gridload(); silence it."""
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def load_png_dir(name, limit=None):
    """Sorted listing of Grids/<name> -> list of float64 grids exactly as gridload makes them
    (Utils/gridmaker.py:87-93: clip(img-1,-1,1)), but in a deterministic (sorted) order."""
    import numpy as np
    from PIL import Image
    d = os.path.join(REFERENCE_ROOT, "Grids", name)
    out = []
    for fname in sorted(os.listdir(d)):
        if limit is not None and len(out) >= limit:
            break
        img = np.array(Image.open(os.path.join(d, fname))).astype(float)
        out.append(np.clip(img - 1, -1, 1))
    return out
