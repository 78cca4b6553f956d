"""Minimal stand-in for `gymnasium` (TEST INFRASTRUCTURE ONLY).

The reference only needs `gym.Env` as a base class, `spaces.Box` as a shape/dtype holder
(`sb3_envs/BaseRLAviary.py:133,267`) and `envs.registration.register` at package import
(`gym_pybullet_drones/__init__.py:1-21`).  gymnasium is not installed in this container.
"""
from . import spaces  # noqa: F401


class Env:
    metadata = {}

    def reset(self, seed=None, options=None):
        raise NotImplementedError

    def step(self, action):
        raise NotImplementedError

    def close(self):
        pass
