"""Harness that runs the UNMODIFIED reference (`/root/reference/gym_pybullet_drones`) headless.

TEST INFRASTRUCTURE ONLY.  It is used to (a) generate the golden fixtures under `tests/golden/`, (b)
cross-check `oracle/cattle_oracle.py` directly against the reference and (c) time the reference's own CPU
code as the baseline arm of `bench.py`.  The reference tree is taken from `/root/reference` (build container)
or, on the GPU box where that path does not exist, from the byte-for-byte staged copy `oracle/_ref`
(`oracle/make_ref.py`, git-ignored, travels like libchs.so).

How the reference is made importable:
  * `oracle/ref_harness/standins/` provides `pybullet`, `pybullet_data`, `gymnasium`
    (not installed here); the reference's own .py files are imported as they are.
  * drone count is pinned through `MIN_NUM_DRONES = MAX_NUM_DRONES = N`
    (`sb3_envs/BaseAviary.py:242,307` draw it with `random.randint`).
  * for M > 16 cows `yaml.safe_load` is patched during construction so that
    `cattle_spawn_data` (`sb3_envs/BaseAviary.py:93-94`) is a synthetic table.
  * a curriculum start level other than the hard-coded 7 (`sb3_envs/CattleAviary.py:62`) is
    obtained by swapping the `CurriculumLearning` symbol seen by that module while constructing.
"""
import contextlib
import io
import os
import random
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_STANDINS = os.path.join(_HERE, "standins")
_STAGED = os.path.abspath(os.path.join(_HERE, "..", "_ref"))


def _find_root():
    for cand in (os.environ.get("CHS_REFERENCE_ROOT"), "/root/reference", _STAGED):
        if cand and os.path.isdir(os.path.join(cand, "gym_pybullet_drones")):
            return cand
    return "/root/reference"


REFERENCE_ROOT = _find_root()


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "gym_pybullet_drones"))


def _ensure_paths():
    # stand-ins go to the END of sys.path: a real pybullet / gymnasium, where one is installed, wins
    if _STANDINS not in sys.path:
        sys.path.append(_STANDINS)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)


def third_party_kinds():
    """Which pybullet / gymnasium the reference is running on: "stand-in" (oracle/ref_harness/standins) or
    "real <version>".  Recorded next to every CPU number."""
    _ensure_paths()
    out = {}
    for name in ("pybullet", "gymnasium"):
        try:
            mod = __import__(name)
            f = os.path.abspath(getattr(mod, "__file__", "") or "")
            out[name] = "stand-in" if f.startswith(_STANDINS) else "real " + str(getattr(mod, "__version__", "?"))
        except ImportError:
            out[name] = "missing"
    return out


def import_reference():
    """Returns (CattleAviary, Physics, ActionType, CurriculumLearning, modules dict)."""
    _ensure_paths()
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with contextlib.redirect_stdout(io.StringIO()):
            from gym_pybullet_drones.sb3_envs import CattleAviary as _pkg  # noqa: F401
            # the package attribute `sb3_envs.CattleAviary` is the class, not the module
            ca_mod = sys.modules["gym_pybullet_drones.sb3_envs.CattleAviary"]
            import gym_pybullet_drones.utils.curriculum_learning as cl_mod
            from gym_pybullet_drones.utils.enums import Physics, ActionType
    return ca_mod.CattleAviary, Physics, ActionType, cl_mod.CurriculumLearning, {"ca": ca_mod, "cl": cl_mod}


def synthetic_spawn_table(n_sims, n_cows, seed=0, min_radius=8.0, max_radius=13.0,
                          min_cow_distance=0.8):
    """Spawn-table generator following the algorithm of `utils/cattle_spawn.py:4-46`.

    The offset half-range is widened from 2 m to max(2, 0.5*sqrt(n_cows)) m so that the 0.8 m
    minimum spacing stays feasible for large herds.
    Returns float64 array (n_sims, n_cows, 2), coordinates rounded to 3 decimals like the YAML.
    """
    rng = np.random.RandomState(seed)
    half = max(2.0, 0.5 * np.sqrt(n_cows))
    out = np.zeros((n_sims, n_cows, 2))
    for s in range(n_sims):
        theta = rng.uniform(0, 2 * np.pi)
        r = rng.uniform(min_radius, max_radius)
        centre = np.array([r * np.cos(theta), r * np.sin(theta)])
        placed = []
        for c in range(n_cows):
            while True:
                off = rng.uniform(-half, half, size=2)
                off *= rng.choice([-1, 1], size=2)
                pos = centre + off
                if all(np.linalg.norm(pos - q) >= min_cow_distance for q in placed):
                    placed.append(pos)
                    out[s, c] = np.round(pos, 3)
                    break
    return out


def table_to_yaml_dict(table):
    return {"simulations": [{"id": s + 1,
                             "cows": [{"id": c + 1, "x": float(table[s, c, 0]), "y": float(table[s, c, 1])}
                                      for c in range(table.shape[1])]}
                            for s in range(table.shape[0])]}


def shipped_spawn_table():
    """The reference's own `config/cattle_positions.yaml` as an array (100, 16, 2)."""
    import yaml
    path = os.path.join(REFERENCE_ROOT, "gym_pybullet_drones", "config", "cattle_positions.yaml")
    with open(path) as f:
        d = yaml.safe_load(f)
    return np.array([[[c["x"], c["y"]] for c in sim["cows"]] for sim in d["simulations"]], dtype=np.float64)


def make_reference_env(n_drones, n_cattle, ctrl_freq=30, pyb_freq=240, level=7, seed=0,
                       spawn_table=None, act="vel"):
    """Constructs the reference `CattleAviary` under Physics.DYN with pinned sizes."""
    CattleAviary, Physics, ActionType, CurriculumLearning, mods = import_reference()
    import yaml
    random.seed(seed)
    np.random.seed(seed)
    orig_load = yaml.safe_load
    orig_cl = mods["ca"].CurriculumLearning
    if spawn_table is not None:
        d = table_to_yaml_dict(np.asarray(spawn_table))
        yaml.safe_load = lambda f: d
    if level != 7:
        mods["ca"].CurriculumLearning = lambda _lvl: orig_cl(level)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            env = CattleAviary(num_drones=12, num_cattle=n_cattle, physics=Physics.DYN,
                               pyb_freq=pyb_freq, ctrl_freq=ctrl_freq,
                               act={"vel": ActionType.VEL, "rpm": ActionType.RPM, "pid": ActionType.PID,
                                    "one_d_rpm": ActionType.ONE_D_RPM, "one_d_pid": ActionType.ONE_D_PID}[act])
    finally:
        yaml.safe_load = orig_load
        mods["ca"].CurriculumLearning = orig_cl
    env.MIN_NUM_DRONES = env.MAX_NUM_DRONES = n_drones
    if act != "vel":
        # `SPEED_LIMIT` only exists for ActionType.VEL (BaseRLAviary.py:101-102) but the reward reads it
        # unconditionally (CattleAviary.py:253): without this the reference raises AttributeError on
        # the first step of every other action type.  Harness-side patch, documented in DESIGN.md.
        env.SPEED_LIMIT = 0.3 * env.MAX_SPEED_KMH * (1000 / 3600)
    return env


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


# ---- state capture --------------------------------------------------------------------------------
NMAX = 12


def _pyb():
    _ensure_paths()
    import pybullet
    return pybullet


class _ZeroCtrl:
    integral_pos_e = last_rpy = integral_rpy_e = np.zeros(3)


def capture_state(env):
    """Everything that `step()` reads, as plain float64/int arrays padded to NMAX drone slots."""
    n = env.NUM_DRONES
    m = env.NUM_CATTLE

    def pad(a, w):
        out = np.zeros((NMAX, w))
        out[:n] = np.asarray(a, dtype=np.float64)[:n]
        return out

    ctrl = getattr(env, "ctrl", None) or [_ZeroCtrl] * NMAX   # ActionType.RPM builds no controllers
    st = {
        "n_drones": np.int32(n),
        "pos": pad(env.pos, 3), "quat": pad(env.quat, 4), "rpy": pad(env.rpy, 3),
        "vel": pad(env.vel, 3), "ang_v": pad(env.ang_v, 3), "rpy_rates": pad(env.rpy_rates, 3),
        "last_rpm": pad(env.last_clipped_action, 4),
        # the controllers are per slot (12 of them) and survive resets
        "pid_ipos": np.array([c.integral_pos_e for c in ctrl], dtype=np.float64).reshape(NMAX, 3),
        "pid_last_rpy": np.array([c.last_rpy for c in ctrl], dtype=np.float64).reshape(NMAX, 3),
        "pid_irpy": np.array([c.integral_rpy_e for c in ctrl], dtype=np.float64).reshape(NMAX, 3),
        "cattle_pos": np.asarray(env.cattle_pos, dtype=np.float64).copy().reshape(m, 3),
        # `_flockingStep` writes the new velocities into the body store (BaseAviary.py:1398-1400);
        # the cached `env.cattle_vel` only catches up at the next kinematics refresh (:425), before
        # anything reads it, so the store is the true state.
        "cattle_vel": np.array([_pyb().getBaseVelocity(cid, physicsClientId=env.CLIENT)[0]
                                for cid in env.CATTLE_IDS], dtype=np.float64).reshape(m, 3),
        "step_counter": np.int64(env.step_counter),
        "step_counter_A": np.int64(env.step_counter_A),
        "prev_cent_dist": np.float64(np.nan if env.prev_cent_dists is None else env.prev_cent_dists),
        "spacing_clock": np.float64(env.drone_spacing_clock),
        "level": np.int32(env.curriculum.curriculum_level),
        "tally": np.int32(env.curriculum.curriculum_success_tally),
        "spawn_idx": np.int32(env.Cattle_Spawn_Index),
    }
    return st


def record_rollout(env, actions, reset_first=True):
    """Steps the reference with the given action sequence, resetting when an episode ends.

    Returns dict of stacked arrays: `before_*` (state entering step t), `after_*` (state leaving
    step t, before any reset), `action`, `obs`, `reward`, `terminated`, `truncated`, and
    `reset_after` (1 where the harness called reset() after step t).
    """
    if reset_first:
        quiet(env.reset)
    rows = []
    n_act = actions.shape[1]
    for t in range(actions.shape[0]):
        before = capture_state(env)
        a = actions[t].astype(np.float32)
        obs, rew, term, trunc, _ = quiet(env.step, a[: max(env.NUM_DRONES, n_act)])
        after = capture_state(env)
        row = {"action": a, "obs": np.asarray(obs, dtype=np.float32), "reward": np.float64(rew),
               "terminated": np.uint8(bool(term)), "truncated": np.uint8(bool(trunc))}
        row.update({"before_" + k: v for k, v in before.items()})
        row.update({"after_" + k: v for k, v in after.items()})
        done = bool(term) or bool(trunc)
        row["reset_after"] = np.uint8(done)
        if done:
            quiet(env.reset)
        rows.append(row)
    return {k: np.stack([r[k] for r in rows]) for k in rows[0]}


# ---- MARL (rllib_envs) variant -----------------------------------------------------------------------
def import_reference_marl():
    """Returns (MARLCattleAviary, Physics, ActionType, module) of the reference's rllib_envs stack."""
    _ensure_paths()
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with contextlib.redirect_stdout(io.StringIO()):
            import gym_pybullet_drones.rllib_envs.MARLCattleAviary as m_mod
            from gym_pybullet_drones.utils.enums import Physics, ActionType
    return m_mod.MARLCattleAviary, Physics, ActionType, m_mod


def make_reference_marl_env(n_drones, n_cattle, ctrl_freq=30, pyb_freq=240, level=0, seed=0, spawn_table=None):
    """Constructs the reference `MARLCattleAviary` (rllib_envs/MARLCattleAviary.py) under Physics.DYN."""
    MARL, Physics, ActionType, m_mod = import_reference_marl()
    import yaml
    random.seed(seed)
    np.random.seed(seed)
    orig_load = yaml.safe_load
    orig_cl = m_mod.CurriculumLearning
    if spawn_table is not None:
        d = table_to_yaml_dict(np.asarray(spawn_table))
        yaml.safe_load = lambda f: d
    if level != 0:
        m_mod.CurriculumLearning = lambda _lvl: orig_cl(level)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            env = MARL(num_drones=12, num_cattle=n_cattle, physics=Physics.DYN, pyb_freq=pyb_freq, ctrl_freq=ctrl_freq,
                       act=ActionType.VEL)
    finally:
        yaml.safe_load = orig_load
        m_mod.CurriculumLearning = orig_cl
    env.MIN_NUM_DRONES = env.MAX_NUM_DRONES = n_drones
    return env


def record_marl_rollout(env, actions, reset_first=True):
    """Like record_rollout for the per-drone dict API (rllib_envs/BaseAviary.py:320-438): obs (12, L), reward,
    terminated and truncated are (12,) arrays padded with zeros beyond NUM_DRONES.  An episode is restarted when
    every drone is terminated or every drone is truncated (`__all__`), or when any flag is set (what a
    VecEnv-style caller does) -- `reset_after` records which."""
    if reset_first:
        quiet(env.reset)
    rows = []
    for t in range(actions.shape[0]):
        before = capture_state(env)
        n = env.NUM_DRONES
        a = actions[t].astype(np.float32)
        obs, rew, term, trunc, _ = quiet(env.step, a[: max(n, 1)])
        after = capture_state(env)
        L = len(obs[0])
        O = np.zeros((NMAX, L), dtype=np.float32)
        R = np.zeros((NMAX,), dtype=np.float64)
        T = np.zeros((NMAX,), dtype=np.uint8)
        U = np.zeros((NMAX,), dtype=np.uint8)
        for i in range(n):
            O[i], R[i], T[i], U[i] = obs[i], rew[i], bool(term[i]), bool(trunc[i])
        row = {"action": a, "obs": O, "reward": R, "terminated": T, "truncated": U,
               "term_all": np.uint8(bool(term["__all__"])), "trunc_all": np.uint8(bool(trunc["__all__"]))}
        row.update({"before_" + k: v for k, v in before.items()})
        row.update({"after_" + k: v for k, v in after.items()})
        done = bool(T[:n].any() or U[:n].any())
        row["reset_after"] = np.uint8(done)
        if done:
            quiet(env.reset)
        rows.append(row)
    return {k: np.stack([r[k] for r in rows]) for k in rows[0]}


# ---- RLlib wrapper (rllib_envs/marl_wrapper.py) ---------------------------------------------------------
def make_reference_wrapper(n_drones, n_cattle, ctrl_freq=30, pyb_freq=240, level=0, seed=0, spawn_table=None):
    """Constructs the reference's `RLlibMultiAgentWrapper` through its own `__init__` (which builds the
    `MARLCattleAviary` from `env_config`), with `ray.rllib.env.MultiAgentEnv` taken from the stand-in package and the
    same harness-side pinning as `make_reference_marl_env` (drone count, spawn table, start level)."""
    MARL, Physics, ActionType, m_mod = import_reference_marl()
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with contextlib.redirect_stdout(io.StringIO()):
            import gym_pybullet_drones.rllib_envs.marl_wrapper as w_mod
    import yaml
    random.seed(seed)
    np.random.seed(seed)
    orig_load = yaml.safe_load
    orig_cl = m_mod.CurriculumLearning
    if spawn_table is not None:
        d = table_to_yaml_dict(np.asarray(spawn_table))
        yaml.safe_load = lambda f: d
    if level != 0:
        m_mod.CurriculumLearning = lambda _lvl: orig_cl(level)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            w = w_mod.RLlibMultiAgentWrapper(dict(num_drones=12, num_cattle=n_cattle, physics=Physics.DYN, pyb_freq=pyb_freq,
                                                  ctrl_freq=ctrl_freq, act=ActionType.VEL))
    finally:
        yaml.safe_load = orig_load
        m_mod.CurriculumLearning = orig_cl
    w.env.MIN_NUM_DRONES = w.env.MAX_NUM_DRONES = n_drones
    return w


def _agent_mask(w):
    return np.int32(sum(1 << int(a.split("_")[1]) for a in w.agents))


def record_wrapper_rollout(w, actions, reset_first=True):
    """Steps the reference WRAPPER (marl_wrapper.py:77-119) with `agent_<i>` action dicts for the agents still in its
    `agents` list.  Outputs are (12,) arrays with zeros for agents that got no entry; `before_/after_active_agents` is the
    wrapper's `agents` list as a bit mask.  An episode is restarted when the wrapper reports "__all__" or -- because its
    "__all__" only ever follows terminations (:116-117) -- when any agent is truncated (`reset_after`)."""
    if reset_first:
        quiet(w.reset)
    rows = []
    for t in range(actions.shape[0]):
        env = w.env
        before = capture_state(env)
        before["active_agents"] = _agent_mask(w)
        n = env.NUM_DRONES
        a = actions[t].astype(np.float32)
        obs, rew, term, trunc, _ = quiet(w.step, {aid: a[int(aid.split("_")[1])] for aid in w.agents})
        after = capture_state(env)
        after["active_agents"] = _agent_mask(w)
        L = env.observation_space.shape[0]
        O = np.zeros((NMAX, L), dtype=np.float32)
        R = np.zeros((NMAX,), dtype=np.float64)
        T = np.zeros((NMAX,), dtype=np.uint8)
        U = np.zeros((NMAX,), dtype=np.uint8)
        for aid in obs:
            i = int(aid.split("_")[1])
            O[i], R[i], T[i], U[i] = obs[aid], rew[aid], bool(term[aid]), bool(trunc[aid])
        row = {"action": a, "obs": O, "reward": R, "terminated": T, "truncated": U,
               "term_all": np.uint8(bool(term["__all__"])), "trunc_all": np.uint8(bool(trunc["__all__"]))}
        row.update({"before_" + k: v for k, v in before.items()})
        row.update({"after_" + k: v for k, v in after.items()})
        done = bool(term["__all__"]) or bool(U[:n].any())
        row["reset_after"] = np.uint8(done)
        if done:
            quiet(w.reset)
        rows.append(row)
    return {k: np.stack([r[k] for r in rows]) for k in rows[0]}
