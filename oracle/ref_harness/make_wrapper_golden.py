"""Generates `tests/golden/wrapper/w_*.npz`: rollouts recorded THROUGH the reference's RLlib wrapper
(`rllib_envs/marl_wrapper.py`, unmodified; `ray.rllib.env.MultiAgentEnv` from the stand-in package).  Build container only,
TEST INFRASTRUCTURE ONLY.

    python oracle/ref_harness/make_wrapper_golden.py

What RLlib sees differs from what `env.step` returned: the wrapper calls the stateful per-agent functions a second time
(:104-110).  These fixtures pin `ChsParams.marl_wrapper_quirks` (the opt-in bug-compatibility mode) the same way the
`m_*` fixtures pin the plain MARL packing: full state before / after every step, the action, and the wrapper's outputs.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_env  # noqa: E402
from make_golden import gentle_actions, ring_around_herd, teleport_drones, uniform_actions, zero_actions  # noqa: E402

OUT = os.path.abspath(os.path.join(HERE, "..", "..", "tests", "golden", "wrapper"))


def _run(name, level, n, m, steps, prep=None, tally=None, clock=None, zero=False, gentle=True, table=None, seed=171):
    table = ref_env.shipped_spawn_table() if table is None else table
    w = ref_env.make_reference_wrapper(n, m, ctrl_freq=30, level=level, seed=seed + level,
                                       spawn_table=None if m <= 16 and table.shape[1] == 16 else table)
    ref_env.quiet(w.reset)
    env = w.env
    if prep is not None:
        prep(env)
    if tally is not None:
        env.curriculum.curriculum_success_tally = tally
    if clock is not None:
        env.drone_spacing_clock = clock
    acts = zero_actions(steps) if zero else (gentle_actions(173 + level, steps) if gentle else uniform_actions(179 + level, steps))
    rec = ref_env.record_wrapper_rollout(w, acts, reset_first=False)
    meta = dict(name=name, n_drones=n, start_level=level, act="vel", packing="marl", wrapper=True, n_cattle=int(env.NUM_CATTLE),
                ctrl_freq=int(env.CTRL_FREQ), pyb_freq=int(env.PYB_FREQ), obs_len=int(rec["obs"].shape[-1]),
                episode_len_sec=float(env.EPISODE_LEN_SEC))
    rec["meta_json"] = np.array(json.dumps(meta))
    rec["spawn_table"] = np.asarray(table, dtype=np.float64)
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **rec)
    print(f"{name}: {steps} steps, resets={int(rec['reset_after'].sum())}, agent-terminations={int(rec['terminated'].sum())}, "
          f"agent-truncations={int(rec['truncated'].sum())}, all={int(rec['term_all'].sum())}, "
          f"masks seen={sorted(set(rec['before_active_agents'].tolist()))}, {os.path.getsize(path) / 1024:.0f} KiB")


def main():
    table20 = ref_env.synthetic_spawn_table(100, 20, seed=0)
    # BASELINE shape, uniform actions: truncations, approach term always zero at the wrapper
    _run("w_3d20c_uniform", 0, 3, 20, 90, gentle=False, table=table20)
    # level 0, clock just below the hold timer, tally just below the threshold: the wrapper's extra calls cross the timer
    # inside the second pass, agents leave the list one after the other, "__all__" fires
    _run("w_l0_hold_levelup", 0, 3, 16, 20, prep=lambda e: teleport_drones(e, [(0, 0), (0.8, 0), (1.6, 0)]), tally=97,
         clock=10 - 5.5 / 30, zero=True)
    _run("w_l2_approach", 2, 4, 16, 30, prep=lambda e: teleport_drones(e, ring_around_herd(e, 1.3, 0.2)), tally=96)
    _run("w_l4_surround", 4, 4, 16, 30, prep=lambda e: teleport_drones(e, ring_around_herd(e, 2.2, 0.4)))


if __name__ == "__main__":
    main()
