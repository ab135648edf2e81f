"""Action spaces — Action_Spaces/{base_space,discrete}.py of the reference."""
import numpy as np


class Base_Space(object):
    def __init__(self):
        super().__init__()

    def sample(self):
        raise NotImplementedError()


class Discrete(Base_Space):
    """``num_actions`` = 4**N for SuperGridRL, 4 for DecGridRL (discrete.py:4-14)."""

    def __init__(self, num_actions):
        super().__init__()
        self.num_actions = num_actions

    def sample(self, s=None):
        return np.random.randint(self.num_actions, size=s)

    def sample_batch(self, num_envs, num_robots, generator=None, device="cuda:0"):
        """Batched extension: uint8 [E, N] uniform per-robot moves drawn on the device."""
        import torch
        return torch.randint(0, 4, (num_envs, num_robots), dtype=torch.uint8, device=device,
                             generator=generator)
