"""`GpuCattleVecEnv`: the batched simulator behind the Stable-Baselines3 `VecEnv` contract.

Drop-in for what `simulator/CTDECattleHerder.py:91-97` builds with
`make_vec_env(CattleAviary, n_envs=24, vec_env_cls=SubprocVecEnv)`: same observation / action spaces per env
(`(12, L)` float32 observation, `(num_drones, 4)` float32 action in [-1, 1]), `reset() -> obs[E,12,L]`,
`step_async(actions)`, `step_wait() -> (obs, rewards, dones, infos)` with SB3's auto-reset semantics:
`dones = terminated | truncated`, `infos[i]["terminal_observation"]` for finished envs and
`infos[i]["TimeLimit.truncated"] = truncated and not terminated`.  numpy arrays in / out (SB3 is numpy based);
host<->device copies go through pinned buffers inside `chs_step_host`.

If stable_baselines3 / gymnasium are importable the class derives from SB3's `VecEnv` and uses gymnasium
spaces; otherwise it derives from a minimal local base with the same surface (neither package is installed in
the build container).
"""
import numpy as np

from .batched_env import CattleHerdingBatchedEnv
from .enums import ActionType, ObservationType, Physics

try:  # pragma: no cover - depends on the box
    from gymnasium import spaces as _spaces
    _Box = _spaces.Box
except Exception:  # noqa: BLE001
    class _Box:
        """Minimal stand-in for gymnasium.spaces.Box (shape / dtype / bounds / sample / contains)."""

        def __init__(self, low, high, shape=None, dtype=np.float32):
            shape = tuple(shape if shape is not None else np.shape(low))
            self.low = np.broadcast_to(np.asarray(low, dtype=dtype), shape).copy()
            self.high = np.broadcast_to(np.asarray(high, dtype=dtype), shape).copy()
            self.shape, self.dtype = shape, np.dtype(dtype)
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)

        def sample(self):
            lo = np.where(np.isfinite(self.low), self.low, -1.0)
            hi = np.where(np.isfinite(self.high), self.high, 1.0)
            return self._rng.uniform(lo, hi).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

try:  # pragma: no cover - depends on the box
    from stable_baselines3.common.vec_env import VecEnv as _VecEnvBase
    HAVE_SB3 = True
except Exception:  # noqa: BLE001
    HAVE_SB3 = False

    class _VecEnvBase:
        def __init__(self, num_envs, observation_space, action_space):
            self.num_envs, self.observation_space, self.action_space = num_envs, observation_space, action_space
            self.render_mode = None

        def step(self, actions):
            self.step_async(actions)
            return self.step_wait()


def observation_space(obs_rows, obs_len):
    """BaseRLAviary._observationSpace (sb3_envs/BaseRLAviary.py:243-267): unbounded float32 (R, L)."""
    return _Box(low=-np.inf, high=np.inf, shape=(obs_rows, obs_len), dtype=np.float32)


def action_space(num_drones, width):
    """BaseRLAviary._actionSpace (:106-133): float32 (num_drones, A) in [-1, 1]."""
    return _Box(low=-1.0, high=1.0, shape=(num_drones, width), dtype=np.float32)


class GpuCattleVecEnv(_VecEnvBase):
    def __init__(self, num_envs, num_drones=12, num_cattle=16, physics=Physics.DYN, pyb_freq=240, ctrl_freq=60,
                 obs=ObservationType.COKIN, act=ActionType.VEL, gui=False, record=False, device=None, seed=0,
                 spawn_table=None, autotune_launch=False, **sim_kwargs):
        """kwargs follow `CattleAviary.__init__` (sb3_envs/CattleAviary.py:14-28) with its defaults for
        frequencies; `physics` must be Physics.DYN (Bullet's PYB solver is not reproducible and not provided);
        `gui` / `record` are refused: there is no graphics stack.  `autotune_launch=True` times the launch shapes of
        `CattleHerdingBatchedEnv.LAUNCH_CANDIDATES` on a probe batch once and keeps the fastest (results are unaffected)."""
        if physics != Physics.DYN:
            raise ValueError("only Physics.DYN is implemented (the PYB constraint solver is not available)")
        if gui or record:
            raise ValueError("gui/record are unavailable (headless simulator)")
        if obs != ObservationType.COKIN:
            raise ValueError("the fork implements ObservationType.COKIN only (BaseRLAviary.py:251-253)")
        if spawn_table is None:
            from . import spawn
            spawn_table = spawn.generate(100, num_cattle, seed=0)
        self.sim = CattleHerdingBatchedEnv(num_envs, num_drones=num_drones, num_cattle=num_cattle, pyb_freq=pyb_freq,
                                           ctrl_freq=ctrl_freq, act=act, auto_reset=True, spawn_table=spawn_table,
                                           device=device, seed=seed, **sim_kwargs)
        if autotune_launch:
            self.sim.autotune_launch()
        super().__init__(num_envs, observation_space(self.sim.obs_rows, self.sim.obs_len),
                         action_space(self.sim.max_drones, self.sim.action_width))
        self.CTRL_FREQ, self.PYB_FREQ = int(ctrl_freq), int(pyb_freq)
        self.CTRL_TIMESTEP = 1.0 / ctrl_freq
        self._actions = None
        self._seed = seed
        self._shared_info = {"answer": 42}     # shared by the envs that did not finish a step: treat as read-only

    # ---- VecEnv contract --------------------------------------------------------------------------------
    def reset(self):
        return self.sim.reset_host().copy()

    def step_async(self, actions):
        self._actions = np.asarray(actions, dtype=np.float32)

    def step_wait(self):
        obs, rew, term, trunc, done_idx, term_obs = self.sim.step_host(self._actions)
        dones = (term | trunc).astype(bool)
        # E Python dicts per step would dominate a 4096-env rollout: envs that did not finish share ONE info dict
        # (`CattleAviary._computeInfo` returns the constant {"answer": 42}, :557-568; SB3 only reads infos of finished
        # envs and copies before it writes), finished envs get their own
        infos = [self._shared_info] * self.num_envs
        for r, e in enumerate(done_idx):
            infos[e] = {"answer": 42, "terminal_observation": term_obs[r].copy(),
                        "TimeLimit.truncated": bool(trunc[e]) and not bool(term[e])}
        return obs.copy(), rew.copy(), dones, infos

    def close(self):
        self.sim.close()

    def seed(self, seed=None):
        """The Philox key is fixed at construction; the reference's `reset(seed=...)` ignores its seed too
        (sb3_envs/BaseAviary.py:280-331)."""
        return [self._seed + i for i in range(self.num_envs)]

    def get_attr(self, attr_name, indices=None):
        n = self.num_envs if indices is None else len(self._indices(indices))
        return [getattr(self, attr_name)] * n

    def set_attr(self, attr_name, value, indices=None):
        setattr(self, attr_name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        n = self.num_envs if indices is None else len(self._indices(indices))
        return [getattr(self, method_name)(*args, **kwargs)] * n

    def env_is_wrapped(self, wrapper_class, indices=None):
        n = self.num_envs if indices is None else len(self._indices(indices))
        return [False] * n

    def get_images(self):
        return [None] * self.num_envs

    def _indices(self, indices):
        if indices is None:
            return list(range(self.num_envs))
        return [indices] if isinstance(indices, int) else list(indices)

    # ---- telemetry ----------------------------------------------------------------------------------------
    def episode_stats(self, reset=True):
        """Device-side episode statistics (dict), the vector that `dist.reduce_stats` all-reduces over NCCL."""
        from ._lib import STAT_NAMES
        s = self.sim.stats(reset=reset)
        return dict(zip(STAT_NAMES, s[: len(STAT_NAMES)].tolist()))
