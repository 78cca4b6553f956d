"""Single-environment facades with the reference's class names and gymnasium contract.

`CattleAviary(BaseRLAviary(BaseAviary))` mirrors `gym_pybullet_drones/sb3_envs/{BaseAviary,BaseRLAviary,
CattleAviary}.py` for `num_envs = 1`: same constructor kwargs, `reset(seed, options) -> (obs, info)`,
`step(action) -> (obs, reward, terminated, truncated, info)`, `action_space`, `observation_space`, and the
attributes the drivers touch (`NUM_DRONES`, `CTRL_FREQ`, `CTRL_TIMESTEP`, `EPISODE_LEN_SEC`, `is_evaluating`,
`evaluation_save()`; simulator/CTDECattleHerder.py:182-218).  All arithmetic runs in the CUDA library; these
classes only pack arguments.  Differences from the reference, all forced by the platform:
  * `physics` defaults to Physics.DYN and no other value is accepted (Bullet's PYB solver is not available);
  * `gui=True` / `record=True` raise (no graphics stack); `render()` prints the text summary only;
  * `num_cattle` > 16 works (a generated spawn table replaces the shipped 16-cow YAML);
  * randomness comes from Philox keyed by `seed` instead of the process-global Mersenne Twister.
"""
import numpy as np
import torch

from .batched_env import CattleHerdingBatchedEnv
from .enums import ActionType, DroneModel, ObservationType, Physics
from .vec_env import action_space as _action_space, observation_space as _observation_space

# episode length / drone-count range per curriculum level (utils/curriculum_learning.py:10-194)
_LEVEL_EP_LEN = {0: 40, 1: 40, 2: 40, 3: 40, 4: 80, 5: 40, 6: 80, 7: 80}
_LEVEL_DRONES = {0: (3, 3), 1: (4, 4), 2: (4, 4), 3: (4, 4), 4: (4, 4), 5: (4, 4), 6: (4, 12), 7: (4, 12)}


class BaseAviary:
    """sb3_envs/BaseAviary.py: simulation core (construction, reset, step, state getters)."""

    CURRICULUM_START_LEVEL = 7          # sb3_envs/CattleAviary.py:62

    def __init__(self, drone_model=DroneModel.CF2X, num_drones=1, num_cattle=1, neighbourhood_radius=np.inf,
                 initial_xyzs=None, initial_rpys=None, physics=Physics.DYN, pyb_freq=240, ctrl_freq=240, gui=False,
                 record=False, obstacles=False, user_debug_gui=False, output_folder="results", obs=ObservationType.COKIN,
                 act=ActionType.VEL, seed=0, spawn_table=None, device=None, precision="fp32", fixed_num_drones=0,
                 curriculum_level=None):
        if drone_model != DroneModel.CF2X:
            raise ValueError("only DroneModel.CF2X is implemented (the training setup of the reference)")
        if physics != Physics.DYN:
            raise ValueError("only Physics.DYN is implemented (the PYB constraint solver is not available)")
        if gui or record:
            raise ValueError("gui/record are unavailable on this platform (headless simulator)")
        if initial_xyzs is not None or initial_rpys is not None:
            raise ValueError("initial_xyzs/initial_rpys are ignored by the reference's reset (BaseAviary.py:315) "
                             "and not supported; use set_state() on the batched env instead")
        level = self.CURRICULUM_START_LEVEL if curriculum_level is None else int(curriculum_level)
        if spawn_table is None:
            from . import spawn
            spawn_table = spawn.generate(100, num_cattle, seed=0)
        self.sim = CattleHerdingBatchedEnv(1, num_drones=num_drones, num_cattle=num_cattle, pyb_freq=pyb_freq,
                                           ctrl_freq=ctrl_freq, act=act, start_level=level,
                                           fixed_num_drones=fixed_num_drones, auto_reset=False, precision=precision,
                                           spawn_table=spawn_table, device=device, seed=seed)
        self.DRONE_MODEL, self.PHYSICS, self.OBS_TYPE, self.ACT_TYPE = drone_model, physics, obs, act
        self.NUM_CATTLE, self.MAX_SLOTS = int(num_cattle), int(num_drones)
        self.PYB_FREQ, self.CTRL_FREQ = int(pyb_freq), int(ctrl_freq)
        self.PYB_STEPS_PER_CTRL = self.PYB_FREQ // self.CTRL_FREQ
        self.CTRL_TIMESTEP, self.PYB_TIMESTEP = 1.0 / self.CTRL_FREQ, 1.0 / self.PYB_FREQ
        self.EPISODE_LEN_SEC = _LEVEL_EP_LEN[level]
        self.MIN_NUM_DRONES, self.MAX_NUM_DRONES = _LEVEL_DRONES[level]
        self.NUM_DRONES = 0
        self.is_evaluating = False
        self.step_counter = 0
        self._device = self.sim.device
        self._tl_seen, self._last_state = 0.0, None
        from .telemetry import EvaluationLog
        self._eval = EvaluationLog(self.sim, 0)
        self.eval_system = self._eval              # the reference's attribute name (BaseAviary.py:124)

    # ---- gymnasium contract -----------------------------------------------------------------------------
    def reset(self, seed=None, options=None):
        """BaseAviary.reset (:280-331).  Like the reference, `seed` does not reseed the environment."""
        obs = self.sim.reset()
        self._refresh()
        self._eval.on_reset(self._last_state)
        return obs[0].cpu().numpy(), self._computeInfo()

    def set_state(self, state):
        """Checkpoint / test hook (no counterpart in the reference): injects a simulator state (the
        fields of `CattleHerdingBatchedEnv.get_state`, leading dimension 1) as if `reset()` had produced it."""
        self.sim.set_state(state)
        self._refresh()
        self._eval.on_reset(self._last_state)

    def step(self, action):
        """BaseAviary.step (:335-465)."""
        a = torch.zeros((1, self.MAX_SLOTS, self.sim.action_width), dtype=torch.float32)
        act = np.asarray(action, dtype=np.float32).reshape(-1, self.sim.action_width)
        a[0, : act.shape[0]] = torch.from_numpy(act[: self.MAX_SLOTS])
        obs, rew, term, trunc = self.sim.step(a.to(self._device))
        out = (obs[0].cpu().numpy(), float(rew[0].item()), bool(term[0].item()), bool(trunc[0].item()), self._computeInfo())
        self._refresh()
        # evaluation_episode_trigger: only an evaluating env, only when the TIME LIMIT is what truncated the episode (the
        # earlier causes return first, CattleAviary.py:509-542), decided on the pre-increment counter like every other time
        # of the log, before this step's sample is appended, and once per _computeTruncated call: _computeReward (:315)
        # and step (BaseAviary.py:460) each make one
        hit = out[3] and self._time_limit_hit()
        if self.is_evaluating and hit:
            self._eval.on_episode_end(self._last_state)
            self._eval.on_episode_end(self._last_state)
        self._eval.on_step(self._last_state)       # update_evaluation_metrics runs every step (BaseAviary.py:462)
        return out

    def _time_limit_hit(self):
        """Did the simulator count a time-limit truncation since the last look?  (`chs_stats` slot 8: the kernel records the
        first cause that fired, in the reference's order.)"""
        count = float(self.sim.stats()[8])
        hit = count > self._tl_seen
        self._tl_seen = count
        return hit

    def render(self, mode="human", close=False):
        """Text summary like BaseAviary.render (:469-494); there is no GUI."""
        st = self.sim.get_state()
        print(f"[INFO] BaseAviary.render() ——— it {int(st['step_counter'][0]):04d} ——— simulation time "
              f"{int(st['step_counter'][0]) * self.PYB_TIMESTEP:.1f}s")
        for i in range(self.NUM_DRONES):
            p, v, r = st["pos"][0, i], st["vel"][0, i], st["rpy"][0, i]
            print(f"[INFO] drone {i} x {p[0]:+06.2f}, y {p[1]:+06.2f}, z {p[2]:+06.2f} ——— velocity {v[0]:+06.2f}, "
                  f"{v[1]:+06.2f}, {v[2]:+06.2f} ——— roll {np.degrees(r[0]):+06.2f}, pitch {np.degrees(r[1]):+06.2f}, "
                  f"yaw {np.degrees(r[2]):+06.2f}")

    def close(self):
        self.sim.close()

    # ---- helpers the reference exposes ---------------------------------------------------------------------
    def getPyBulletClient(self):
        return -1            # there is no Bullet client

    def getDroneIds(self):
        return np.arange(1, self.NUM_DRONES + 1)

    def evaluation_save(self, save_path="evaluation_data.pkl"):
        """`evaluator.save_evaluation_data` (utils/evaluation.py:73-94): pickles the trajectory log in the reference's
        schema (telemetry.EvaluationLog) and returns the device-side episode statistics vector as well."""
        self._eval.save(save_path)
        return self.sim.stats()

    def _getDroneStateVector(self, nth_drone):
        """(20,) = pos quat rpy vel ang_v last_clipped_action (BaseAviary.py:744-766)."""
        s = self.sim.get_state()
        i = int(nth_drone)
        return np.hstack([s["pos"][0, i], s["quat"][0, i], s["rpy"][0, i], s["vel"][0, i], s["ang_v"][0, i], s["last_rpm"][0, i]])

    def _getCowStateVector(self, nth_cow):
        """(16,) = pos quat rpy vel ang_v (BaseAviary.py:770-789); cattle keep identity attitude on the DYN path."""
        s = self.sim.get_state()
        j = int(nth_cow)
        return np.hstack([s["cattle_pos"][0, j], [0, 0, 0, 1], [0, 0, 0], s["cattle_vel"][0, j], [0, 0, 0]])

    def _refresh(self):
        s = self._last_state = self.sim.get_state()
        self.NUM_DRONES = int(s["n_drones"][0])
        self.step_counter = int(s["step_counter"][0])

    def _computeInfo(self):
        return {"answer": 42}    # CattleAviary._computeInfo (:557-568)


class BaseRLAviary(BaseAviary):
    """sb3_envs/BaseRLAviary.py: action / observation spaces."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.ACTION_BUFFER_SIZE = int(0.15 * self.CTRL_FREQ)       # BaseRLAviary.py:69
        self.action_space = _action_space(self.MAX_SLOTS, self.sim.action_width)
        self.observation_space = _observation_space(self.sim.obs_rows, self.sim.obs_len)

    def HerdCentroid(self):
        s = self.sim.get_state()
        c = s["cattle_pos"][0, :, :2].mean(axis=0)
        return np.array([c[0], c[1], 0.45 + 0.5])               # BaseRLAviary.py:348-366

    def DroneCentroid(self):
        s = self.sim.get_state()
        c = s["pos"][0, : self.NUM_DRONES, :2].mean(axis=0)
        return np.array([c[0], c[1], 0.45 + 0.5])               # BaseRLAviary.py:372-390


class CattleAviary(BaseRLAviary):
    """sb3_envs/CattleAviary.py:14-28: same kwargs and defaults, except `physics` (DYN, see module docstring)."""

    def __init__(self, drone_model=DroneModel.CF2X, num_drones=2, num_cattle=1, neighbourhood_radius=np.inf,
                 initial_xyzs=None, initial_rpys=None, physics=Physics.DYN, pyb_freq=240, ctrl_freq=60, gui=False,
                 record=False, obs=ObservationType.COKIN, act=ActionType.VEL, **kwargs):
        super().__init__(drone_model=drone_model, num_drones=num_drones, num_cattle=num_cattle,
                         neighbourhood_radius=neighbourhood_radius, initial_xyzs=initial_xyzs, initial_rpys=initial_rpys,
                         physics=physics, pyb_freq=pyb_freq, ctrl_freq=ctrl_freq, gui=gui, record=record, obs=obs, act=act,
                         **kwargs)
