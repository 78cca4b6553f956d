"""Generates `tests/golden/*.npz` by running the UNMODIFIED reference (build container only).

    python oracle/ref_harness/make_golden.py [name ...]

Each fixture is one recorded rollout of `gym_pybullet_drones.sb3_envs.CattleAviary` under
`Physics.DYN` (see `ref_env.py` for how it is made importable): per step the full state entering
the step, the float32 action, the returned (obs, reward, terminated, truncated) and the full state
leaving the step.  Scenarios cover the BASELINE.json shapes plus the branches the random-action
rollouts never reach (drones inside the herd, every curriculum level's success branch, level-up,
two-row spawn, K=4, other action types).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_env  # noqa: E402

GOLDEN = os.path.abspath(os.path.join(HERE, "..", "..", "tests", "golden"))


def uniform_actions(seed, steps, width=4):
    return np.random.default_rng(seed).uniform(-1, 1, (steps, 12, width)).astype(np.float32)


def gentle_actions(seed, steps):
    a = np.random.default_rng(seed).uniform(-1, 1, (steps, 12, 4)).astype(np.float32)
    a[..., 3] = np.abs(a[..., 3]) * 0.05
    return a


def teleport_drones(env, xy):
    """Moves the reference's drones through its own pybullet handle, then refreshes kinematics."""
    import pybullet as p
    for i in range(env.NUM_DRONES):
        p.resetBasePositionAndOrientation(env.DRONE_IDS[i], [xy[i][0], xy[i][1], 0.45], [0, 0, 0, 1],
                                          physicsClientId=env.CLIENT)
    env._updateAndStoreKinematicInformation()


def ring_around_herd(env, radius, phase=0.0):
    c = np.mean(np.asarray(env.cattle_pos)[:, :2], axis=0)
    n = env.NUM_DRONES
    return [(c[0] + radius * np.cos(phase + 2 * np.pi * i / n), c[1] + radius * np.sin(phase + 2 * np.pi * i / n))
            for i in range(n)]


SCENARIOS = {}


def scenario(fn):
    SCENARIOS[fn.__name__] = fn
    return fn


def _finish(name, env, rec, meta, table):
    meta = dict(meta)
    meta.update(name=name, n_cattle=int(env.NUM_CATTLE), ctrl_freq=int(env.CTRL_FREQ), pyb_freq=int(env.PYB_FREQ),
                obs_len=int(rec["obs"].shape[-1]), episode_len_sec=float(env.EPISODE_LEN_SEC))
    rec["meta_json"] = np.array(json.dumps(meta))
    rec["spawn_table"] = np.asarray(table, dtype=np.float64)
    os.makedirs(GOLDEN, exist_ok=True)
    path = os.path.join(GOLDEN, name + ".npz")
    np.savez_compressed(path, **rec)
    print(f"{name}: {rec['action'].shape[0]} steps, resets={int(rec['reset_after'].sum())}, "
          f"term={int(rec['terminated'].sum())}, trunc={int(rec['truncated'].sum())}, "
          f"{os.path.getsize(path) / 1024:.0f} KiB")


@scenario
def a_3d20c_uniform():
    """BASELINE config 1 shape: 3 drones / 20 cattle, 240/30 Hz, level 7, uniform actions."""
    table = ref_env.synthetic_spawn_table(100, 20, seed=0)
    env = ref_env.make_reference_env(3, 20, ctrl_freq=30, spawn_table=table, seed=0)
    rec = ref_env.record_rollout(env, uniform_actions(1, 300))
    _finish("a_3d20c_uniform", env, rec, dict(n_drones=3, start_level=7, act="vel"), table)


@scenario
def a_3d20c_gentle():
    """Long episodes (slow commands): exercises many flocking updates and the time-limit branch."""
    table = ref_env.synthetic_spawn_table(100, 20, seed=0)
    env = ref_env.make_reference_env(3, 20, ctrl_freq=30, spawn_table=table, seed=3)
    rec = ref_env.record_rollout(env, gentle_actions(2, 360))
    _finish("a_3d20c_gentle", env, rec, dict(n_drones=3, start_level=7, act="vel"), table)


@scenario
def b_3d16c_60hz():
    """The reference's shipped spawn table and default ctrl_freq=60 (K=4, L=86)."""
    table = ref_env.shipped_spawn_table()
    env = ref_env.make_reference_env(3, 16, ctrl_freq=60, seed=5)
    rec = ref_env.record_rollout(env, uniform_actions(7, 200))
    _finish("b_3d16c_60hz", env, rec, dict(n_drones=3, start_level=7, act="vel"), table)


@scenario
def c_8d100c_uniform():
    """BASELINE config 3 shape: 8 drones / 100 cattle."""
    table = ref_env.synthetic_spawn_table(8, 100, seed=0)
    env = ref_env.make_reference_env(8, 100, ctrl_freq=30, spawn_table=table, seed=11)
    rec = ref_env.record_rollout(env, uniform_actions(13, 40))
    _finish("c_8d100c_uniform", env, rec, dict(n_drones=8, start_level=7, act="vel"), table)


@scenario
def d_inherd_4d16c():
    """Drones start on a ring inside/around the herd: delta gradient, predator repulsion,
    non-zero herding effectiveness, nearest-cow reward term."""
    table = ref_env.shipped_spawn_table()
    env = ref_env.make_reference_env(4, 16, ctrl_freq=30, seed=17)
    ref_env.quiet(env.reset)
    teleport_drones(env, ring_around_herd(env, 2.6, 0.3))
    rec1 = ref_env.record_rollout(env, gentle_actions(19, 60), reset_first=False)
    ref_env.quiet(env.reset)
    teleport_drones(env, ring_around_herd(env, 0.9, 1.1))
    rec2 = ref_env.record_rollout(env, gentle_actions(23, 60), reset_first=False)
    rec = {k: np.concatenate([rec1[k], rec2[k]]) for k in rec1}
    _finish("d_inherd_4d16c", env, rec, dict(n_drones=4, start_level=7, act="vel"), table)


@scenario
def e_12d16c_tworow():
    """12 drones: two-row spawn, neighbour ranking among many candidates."""
    table = ref_env.shipped_spawn_table()
    env = ref_env.make_reference_env(12, 16, ctrl_freq=30, seed=29)
    rec = ref_env.record_rollout(env, uniform_actions(31, 60))
    _finish("e_12d16c_tworow", env, rec, dict(n_drones=12, start_level=7, act="vel"), table)


def zero_actions(steps):
    return np.zeros((steps, 12, 4), dtype=np.float32)


def _level_run(name, level, n, prep, steps=40, tally=None, clock=None, m=16, zero=False):
    table = ref_env.shipped_spawn_table()
    env = ref_env.make_reference_env(n, m, ctrl_freq=30, level=level, seed=37 + level)
    ref_env.quiet(env.reset)
    prep(env)
    if tally is not None:
        env.curriculum.curriculum_success_tally = tally
    if clock is not None:
        env.drone_spacing_clock = clock
    acts = zero_actions(steps) if zero else gentle_actions(41 + level, steps)
    rec = ref_env.record_rollout(env, acts, reset_first=False)
    _finish(name, env, rec, dict(n_drones=n, start_level=level, act="vel"), table)


@scenario
def l0_spacing_hold():
    """Level 0: drones 0.8 m apart; spacing clock preset just below the 10 s hold so that the
    success branch (and the second-call quirk) fires; tally preset so the level bumps to 1."""
    _level_run("l0_spacing_hold", 0, 3, lambda e: teleport_drones(e, [(0, 0), (0.8, 0), (1.6, 0)]),
               steps=30, tally=99, clock=10 - 7.5 / 240)


@scenario
def l0_levelup():
    """Level 0 -> 1: the in-reward termination call crosses the hold timer first, bumps the tally to
    100 and the level to 1; the second call of the same step then runs at level 1 (hold 25 s) and
    returns False, so the episode goes on under the new level's weights."""
    _level_run("l0_levelup", 0, 3, lambda e: teleport_drones(e, [(0, 0), (0.8, 0), (1.6, 0)]),
               steps=24, tally=99, clock=10 - 8.5 / 240, zero=True)


@scenario
def l1_complex_reward():
    _level_run("l1_complex_reward", 1, 4, lambda e: teleport_drones(e, [(0, 0), (0.85, 0), (0, 0.9), (2.4, 0.1)]),
               steps=30, clock=25 - 30.5 / 240, zero=True)


@scenario
def l2_approach():
    _level_run("l2_approach", 2, 4, lambda e: teleport_drones(e, ring_around_herd(e, 1.3, 0.2)), steps=30, tally=99)


@scenario
def l3_approach():
    _level_run("l3_approach", 3, 4, lambda e: teleport_drones(e, ring_around_herd(e, 2.0, 0.7)), steps=30)


@scenario
def l4_surround():
    _level_run("l4_surround", 4, 4, lambda e: teleport_drones(e, ring_around_herd(e, 2.2, 0.4)), steps=30)


@scenario
def l5_surround_spacing():
    _level_run("l5_surround_spacing", 5, 4, lambda e: teleport_drones(e, ring_around_herd(e, 0.58, 0.1)), steps=30,
               m=1, zero=True)


@scenario
def l6_surround_var():
    _level_run("l6_surround_var", 6, 6, lambda e: teleport_drones(e, ring_around_herd(e, 2.8, 0.0)), steps=30)


def _act_run(name, act, width, steps=40):
    table = ref_env.shipped_spawn_table()
    env = ref_env.make_reference_env(3, 16, ctrl_freq=30, seed=53, act=act)
    a = uniform_actions(59, steps, width)
    if act == "pid":      # waypoints near the drones
        a = (a * 1.5 + np.array([1.75, 0.0, 0.45], dtype=np.float32)).astype(np.float32)
    rec = ref_env.record_rollout(env, a)
    _finish(name, env, rec, dict(n_drones=3, start_level=7, act=act), table)


@scenario
def act_rpm():
    _act_run("act_rpm", "rpm", 4)


@scenario
def act_pid():
    _act_run("act_pid", "pid", 3)


@scenario
def act_one_d_rpm():
    _act_run("act_one_d_rpm", "one_d_rpm", 1)


@scenario
def act_one_d_pid():
    _act_run("act_one_d_pid", "one_d_pid", 1)


# ---- MARL (rllib_envs) fixtures ------------------------------------------------------------------------
def _marl_run(name, level, n, m, steps, prep=None, tally=None, clock=None, zero=False, gentle=True, table=None, seed=71):
    table = ref_env.shipped_spawn_table() if table is None else table
    env = ref_env.make_reference_marl_env(n, m, ctrl_freq=30, level=level, seed=seed + level,
                                          spawn_table=None if m <= 16 and table.shape[1] == 16 else table)
    ref_env.quiet(env.reset)
    if prep is not None:
        prep(env)
    if tally is not None:
        env.curriculum.curriculum_success_tally = tally
    if clock is not None:
        env.drone_spacing_clock = clock
    acts = zero_actions(steps) if zero else (gentle_actions(73 + level, steps) if gentle else uniform_actions(79 + level, steps))
    rec = ref_env.record_marl_rollout(env, acts, reset_first=False)
    _finish(name, env, rec, dict(n_drones=n, start_level=level, act="vel", packing="marl"), table)


@scenario
def m_3d20c_uniform():
    """MARL packing, the reference's default level 0 (3 drones), BASELINE shape, uniform actions."""
    table = ref_env.synthetic_spawn_table(100, 20, seed=0)
    _marl_run("m_3d20c_uniform", 0, 3, 20, 120, gentle=False, table=table)


@scenario
def m_l0_hold_levelup():
    """Level 0: spacing clock just below the hold timer and tally just below the threshold: the 2N
    _computeTerminated calls of one step cross the timer, per-drone end-of-episode rewards fire and the level
    bumps in the middle of the per-drone reward loop."""
    _marl_run("m_l0_hold_levelup", 0, 3, 16, 20, prep=lambda e: teleport_drones(e, [(0, 0), (0.8, 0), (1.6, 0)]),
              tally=98, clock=10 - 3.5 / 30, zero=True)


@scenario
def m_l2_approach():
    _marl_run("m_l2_approach", 2, 4, 16, 30, prep=lambda e: teleport_drones(e, ring_around_herd(e, 1.3, 0.2)), tally=97)


@scenario
def m_l4_surround():
    _marl_run("m_l4_surround", 4, 4, 16, 30, prep=lambda e: teleport_drones(e, ring_around_herd(e, 2.2, 0.4)))


@scenario
def m_l5_surround_spacing():
    _marl_run("m_l5_surround_spacing", 5, 4, 1, 20, prep=lambda e: teleport_drones(e, ring_around_herd(e, 0.58, 0.1)), zero=True)


@scenario
def m_l7_6d():
    _marl_run("m_l7_6d", 7, 6, 16, 60, gentle=False)


if __name__ == "__main__":
    names = sys.argv[1:] or list(SCENARIOS)
    for nm in names:
        SCENARIOS[nm]()
