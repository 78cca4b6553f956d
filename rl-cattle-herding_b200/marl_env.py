"""Multi-agent (RLlib) front ends over the MARL packing of the CUDA step.

* `MARLCattleAviary`       -- single environment with the dict contract of
  `gym_pybullet_drones/rllib_envs/MARLCattleAviary.py` + `rllib_envs/BaseAviary.py:280-438`:
  `reset() -> ({i: obs_i}, {"__all__": {}})`, `step(action (N, A)) -> obs, reward, terminated, truncated, info`
  dicts keyed by drone index, the flag dicts carrying "__all__".
* `RLlibMultiAgentWrapper` -- the `agent_<i>` keyed wrapper of `rllib_envs/marl_wrapper.py:8-125`
  (`possible_agents`, `agents`, per-agent spaces, finished agents leave `agents`).  The reference wrapper discards
  what `env.step` returned and calls the stateful `_computeReward / _computeTerminated / _computeTruncated` a second
  time (marl_wrapper.py:104-110), which advances the spacing clock and the curriculum tally twice per step and makes
  the approach term always zero.  By default this wrapper hands out the outputs of the step itself;
  `env_config["reference_wrapper_quirks"] = True` reproduces the reference wrapper bug for bug (the second invocation
  runs inside the step kernel, `ChsParams.marl_wrapper_quirks`; pinned by tests/golden/wrapper/, recorded through the
  reference's wrapper).
* `GpuMARLBatchedEnv`      -- E environments at once, arrays in / arrays out (the per-drone dicts as `(E, D)` arrays).

All arithmetic runs in `libchs.so`; there is no CPU path.
"""
import numpy as np
import torch

from .batched_env import CattleHerdingBatchedEnv
from .enums import ActionType, DroneModel, ObservationType, Physics
from .vec_env import _Box

try:  # pragma: no cover - ray is not installed in the build container
    from ray.rllib.env import MultiAgentEnv as _MultiAgentEnv
except Exception:  # noqa: BLE001
    class _MultiAgentEnv:
        def __init__(self):
            pass


def _spaces(sim):
    """Per-agent spaces (rllib_envs/BaseMARLAviary.py:105-130, 228-248): action (A,) in [-1, 1], observation (L,)."""
    return (_Box(low=-1.0, high=1.0, shape=(sim.action_width,), dtype=np.float32),
            _Box(low=-np.inf, high=np.inf, shape=(sim.obs_len,), dtype=np.float32))


class GpuMARLBatchedEnv(CattleHerdingBatchedEnv):
    """`CattleHerdingBatchedEnv` with `packing="marl"` and the MARL class's defaults (curriculum level 0,
    MARLCattleAviary.py:62; ctrl_freq 60, :23)."""

    def __init__(self, num_envs, num_drones=4, num_cattle=16, ctrl_freq=60, start_level=0, **kw):
        kw.pop("packing", None)
        super().__init__(num_envs, num_drones=num_drones, num_cattle=num_cattle, ctrl_freq=ctrl_freq,
                         start_level=start_level, packing="marl", **kw)

    def all_done(self):
        """(E,) bool: either "__all__" flag of the last step (what RLlib ends the episode on)."""
        D = self.max_drones
        return (self.terminated[:, D] | self.truncated[:, D]).bool()


class MARLCattleAviary:
    """Single environment, reference kwargs (rllib_envs/MARLCattleAviary.py:15-28)."""

    def __init__(self, drone_model=DroneModel.CF2X, num_drones=2, num_cattle=1, neighbourhood_radius=np.inf,
                 initial_xyzs=None, initial_rpys=None, physics=Physics.DYN, pyb_freq=240, ctrl_freq=60, gui=False,
                 record=False, obs=ObservationType.COKIN, act=ActionType.VEL, seed=0, spawn_table=None, device=None,
                 precision="fp32", fixed_num_drones=0, curriculum_level=0, wrapper_quirks=False):
        if drone_model != DroneModel.CF2X:
            raise ValueError("only DroneModel.CF2X is implemented")
        if physics != Physics.DYN:
            raise ValueError("only Physics.DYN is implemented (the PYB constraint solver is not available)")
        if gui or record:
            raise ValueError("gui/record are unavailable on this platform (headless simulator)")
        if spawn_table is None:
            from . import spawn
            spawn_table = spawn.generate(100, num_cattle, seed=0)
        self.sim = GpuMARLBatchedEnv(1, num_drones=num_drones, num_cattle=num_cattle, pyb_freq=pyb_freq,
                                     ctrl_freq=ctrl_freq, act=act, start_level=int(curriculum_level),
                                     fixed_num_drones=fixed_num_drones, auto_reset=False, precision=precision,
                                     spawn_table=spawn_table, device=device, seed=seed,
                                     marl_wrapper_quirks=bool(wrapper_quirks))
        self.wrapper_quirks = bool(wrapper_quirks)
        self.MAX_SLOTS, self.NUM_CATTLE = int(num_drones), int(num_cattle)
        self.PYB_FREQ, self.CTRL_FREQ = int(pyb_freq), int(ctrl_freq)
        self.CTRL_TIMESTEP = 1.0 / self.CTRL_FREQ
        self.action_space, self.observation_space = _spaces(self.sim)
        self.NUM_DRONES = 0
        self.step_counter = 0
        self._last = None

    def _refresh(self):
        s = self.sim.get_state()
        self.NUM_DRONES, self.step_counter = int(s["n_drones"][0]), int(s["step_counter"][0])

    def reset(self, seed=None, options=None):
        obs = self.sim.reset()[0].cpu().numpy()
        self._refresh()
        self._last = obs
        return {i: obs[i] for i in range(self.NUM_DRONES)}, {"__all__": {}}

    def step(self, action):
        n, D, A = self.NUM_DRONES, self.MAX_SLOTS, self.sim.action_width
        a = np.zeros((1, D, A), dtype=np.float32)
        if isinstance(action, dict):
            for i, v in action.items():
                a[0, int(i)] = np.asarray(v, dtype=np.float32).reshape(A)
        else:
            act = np.asarray(action, dtype=np.float32).reshape(-1, A)
            a[0, : min(D, act.shape[0])] = act[:D]
        obs, rew, term, trunc = [x[0].cpu().numpy() for x in self.sim.step(torch.from_numpy(a).to(self.sim.device))]
        self._last = obs
        out_obs = {i: obs[i] for i in range(n)}
        reward = {i: float(rew[i]) for i in range(n)}
        done = {i: bool(term[i]) for i in range(n)}
        done["__all__"] = bool(term[D])
        truncated = {i: bool(trunc[i]) for i in range(n)}
        truncated["__all__"] = bool(trunc[D])
        info = {i: {"answer": 42} for i in range(n)}       # MARLCattleAviary._computeInfo (:388-399)
        self._refresh()
        return out_obs, reward, done, truncated, info

    def _computeObs(self, i):
        return self._last[int(i)]

    def render(self, mode="human"):
        s = self.sim.get_state()
        print(f"[INFO] step {self.step_counter} drones {self.NUM_DRONES} level {int(s['level'][0])}")

    def close(self):
        self.sim.close()


class RLlibMultiAgentWrapper(_MultiAgentEnv):
    """`agent_<i>` keyed wrapper (rllib_envs/marl_wrapper.py:8-125); see the module docstring for the one difference."""

    def __init__(self, env_config=None):
        super().__init__()
        cfg = dict(env_config or {})
        cfg.setdefault("gui", False)
        # opt-in: report what the REFERENCE wrapper reports (its second invocation of the stateful per-agent functions)
        cfg["wrapper_quirks"] = bool(cfg.pop("reference_wrapper_quirks", False))
        self.env = MARLCattleAviary(**cfg)
        self._set_agents(self.env.MAX_SLOTS)

    def _set_agents(self, n):
        self.possible_agents = [f"agent_{i}" for i in range(n)]
        self.agents = self.possible_agents.copy()
        self.action_space = {aid: self.env.action_space for aid in self.possible_agents}
        self.observation_space = {aid: self.env.observation_space for aid in self.possible_agents}

    def get_action_space(self, agent_id):
        return self.env.action_space

    def get_observation_space(self, agent_id):
        return self.env.observation_space

    def reset(self, *, seed=None, options=None):
        obs, _ = self.env.reset(seed=seed, options=options)
        self._set_agents(self.env.NUM_DRONES)
        return {f"agent_{i}": o for i, o in obs.items()}, {f"agent_{i}": {} for i in obs}

    def step(self, action_dict):
        A = self.env.sim.action_width
        action = np.zeros((self.env.NUM_DRONES, A), dtype=np.float32)
        for i, aid in enumerate(self.possible_agents):
            if aid in action_dict and aid in self.agents:
                action[i] = np.asarray(action_dict[aid], dtype=np.float32)
        o, r, d, t, info = self.env.step(action)
        obs, rewards, dones, truncs, infos = {}, {}, {}, {}, {}
        for aid in list(self.agents):
            i = int(aid.split("_")[1])
            obs[aid], rewards[aid], dones[aid], truncs[aid], infos[aid] = o[i], r[i], d[i], t[i], info[i]
        self.agents = [aid for aid in self.agents if not dones.get(aid, False)]
        dones["__all__"] = len(self.agents) == 0
        truncs["__all__"] = len(self.agents) == 0
        if self.env.wrapper_quirks:      # the kernel keeps the same list (ChsStateView.active_agents): they must agree
            assert dones["__all__"] == d["__all__"], "agent list out of step with the simulator's"
        return obs, rewards, dones, truncs, infos

    def render(self, mode="human"):
        return self.env.render(mode=mode)

    def close(self):
        return self.env.close()
