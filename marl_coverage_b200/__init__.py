"""marl_coverage_b200 — B200-native batched MARL-Coverage environment step.

Only the env-step hot path of ExistentialRobotics/MARL-Coverage lives here: the
SuperGridRL / DecGridRL reset()/step() path, the square and lidar sensors, the discrete moves and
the map supply feeding them, as sm_100a CUDA kernels behind the C ABI of include/mcov.h."""
from . import _lib                                              # noqa: F401
from .action_spaces import Base_Space, Discrete                 # noqa: F401
from .maps import MapTable, gridgen, gridload, load_grid_set    # noqa: F401
from .sensors import LidarSensor, Sensor, SquareSensor          # noqa: F401
from .vec_env import VecDecGridRL, VecSuperGridRL               # noqa: F401
from .envs import DecGridRL, SuperGridRL                        # noqa: F401
from .replay import ReplayBuffer, collect                       # noqa: F401
from .sim_env import SuperGrid_Sim, VecSuperGridSim             # noqa: F401
from .rollout import RolloutGraph                               # noqa: F401

__all__ = ["DecGridRL", "SuperGridRL", "VecDecGridRL", "VecSuperGridRL", "Sensor", "SquareSensor",
           "LidarSensor", "Discrete", "ReplayBuffer", "collect", "SuperGrid_Sim", "VecSuperGridSim", "Base_Space", "gridload", "gridgen", "load_grid_set", "MapTable", "RolloutGraph"]
