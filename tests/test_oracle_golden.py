"""Pins `oracle/cattle_oracle.py` against the golden vectors recorded from the reference itself.

CPU only.  Teacher-forced: every recorded step is replayed from the reference's own `before` state and
must reproduce the reference's outputs and `after` state to 1e-9 (float64 vs float64; differences are
summation order and the scipy Euler round trip, both ~1e-13) with bit-exact integers and flags.
Free-running: each recorded episode is replayed from its first state without re-injection.
"""
import numpy as np
import pytest

from _golden import FLOAT_STATE, INT_STATE, Golden, golden_names, max_err

TOL = 1e-9


@pytest.mark.parametrize("name", golden_names())
def test_teacher_forced(name):
    g = Golden(name)
    env = g.make_oracle()
    worst = 0.0
    for t in range(g.steps):
        env.set_state(g.state(t, "before"))
        obs, rew, term, trunc, _ = env.step(g.action(t))
        after = g.state(t, "after")
        got = env.get_state()
        for k in INT_STATE:
            assert int(got[k]) == int(after[k]), (name, t, k, got[k], after[k])
        assert term == bool(g.z["terminated"][t]), (name, t, "terminated")
        assert trunc == bool(g.z["truncated"][t]), (name, t, "truncated")
        for k in FLOAT_STATE:
            e = max_err(got[k], after[k])
            scale = max(1.0, float(np.nanmax(np.abs(np.where(np.isfinite(after[k]), after[k], 0.0)))))
            assert e <= TOL * scale, (name, t, k, e)
            worst = max(worst, e / scale)
        # the observation is float32 in the reference: identical float64 inputs must round identically
        # except at float32 rounding boundaries, so allow 1 ulp there
        ref_obs = g.z["obs"][t]
        assert obs.shape == ref_obs.shape and obs.dtype == np.float32
        assert max_err(obs, ref_obs) <= 2e-7 * max(1.0, float(np.max(np.abs(ref_obs)))), (name, t, "obs")
        e = max_err(rew, g.z["reward"][t])
        assert e <= TOL * max(1.0, abs(float(g.z["reward"][t])) if np.isfinite(g.z["reward"][t]) else 1.0), (name, t, "reward", e)
    assert worst <= TOL


@pytest.mark.parametrize("name", ["a_3d20c_uniform", "b_3d16c_60hz", "d_inherd_4d16c", "e_12d16c_tworow"])
def test_free_running_episodes(name):
    """Replays whole episodes without re-injecting state; the chaotic roll loop
    amplifies 1e-16 differences, so the bound is looser and scaled by episode length."""
    g = Golden(name)
    env = g.make_oracle()
    start = 0
    checked = 0
    for t in range(g.steps):
        if t == start or not np.array_equal(g.z["before_pos"][t], g.z["after_pos"][t - 1]):
            env.set_state(g.state(t, "before"))   # episode start or a harness teleport
        obs, rew, term, trunc, _ = env.step(g.action(t))
        assert term == bool(g.z["terminated"][t]) and trunc == bool(g.z["truncated"][t]), (name, t)
        assert max_err(env.pos, g.z["after_pos"][t]) <= 1e-6, (name, t)
        assert max_err(env.cattle_vel, g.z["after_cattle_vel"][t]) <= 1e-6, (name, t)
        if g.z["reset_after"][t]:
            start = t + 1
            checked += 1
    assert checked >= 1


@pytest.mark.parametrize("name", golden_names("marl"))
def test_marl_teacher_forced(name):
    """MARL packing (rllib_envs): per-drone obs / reward / flags, the sequential reward loop with its
    mid-step curriculum bumps, clock += 1/CTRL_FREQ per call, control-step counter."""
    g = Golden(name)
    assert g.meta["packing"] == "marl"
    env = g.make_oracle()
    n_term = n_trunc = 0
    for t in range(g.steps):
        env.set_state(g.state(t, "before"))
        obs, rew, term, trunc = env.step_marl(g.action(t))
        after = g.state(t, "after")
        got = env.get_state()
        for k in INT_STATE:
            assert int(got[k]) == int(after[k]), (name, t, k, got[k], after[k])
        assert np.array_equal(term, g.z["terminated"][t]), (name, t, "terminated")
        assert np.array_equal(trunc, g.z["truncated"][t]), (name, t, "truncated")
        for k in FLOAT_STATE:
            e = max_err(got[k], after[k])
            scale = max(1.0, float(np.nanmax(np.abs(np.where(np.isfinite(after[k]), after[k], 0.0)))))
            assert e <= TOL * scale, (name, t, k, e)
        ref_obs = g.z["obs"][t]
        assert obs.shape == ref_obs.shape and obs.dtype == np.float32
        assert max_err(obs, ref_obs) <= 2e-7 * max(1.0, float(np.max(np.abs(ref_obs)))), (name, t, "obs")
        assert max_err(rew, g.z["reward"][t]) <= TOL * max(1.0, float(np.max(np.abs(g.z["reward"][t])))), (name, t, "reward")
        n_term += int(term.any()); n_trunc += int(trunc.any())
    assert n_term + n_trunc >= 1


@pytest.mark.parametrize("name", ["w_3d20c_uniform", "w_l0_hold_levelup", "w_l2_approach", "w_l4_surround"])
def test_rllib_wrapper_teacher_forced(name):
    """The reference's RLlib wrapper (rllib_envs/marl_wrapper.py:77-119), recorded through the wrapper itself
    (tests/golden/wrapper, oracle/ref_harness/make_wrapper_golden.py): second invocation of the stateful per-agent
    functions for the agents in its list, list update, "__all__" = list empty."""
    import json, os
    from _golden import GOLDEN_DIR
    z = np.load(os.path.join(GOLDEN_DIR, "wrapper", name + ".npz"))
    meta = json.loads(str(z["meta_json"]))
    import cattle_oracle as co
    env = co.OracleEnv(meta["n_cattle"], ctrl_freq=meta["ctrl_freq"], pyb_freq=meta["pyb_freq"], start_level=meta["start_level"],
                       spawn_table=z["spawn_table"], act=meta["act"])
    masks = set()
    for t in range(z["action"].shape[0]):
        env.set_state({k: z[f"before_{k}"][t] for k in co.STATE_FIELDS})
        obs, rew, term, trunc, mask = env.step_rllib_wrapper(z["action"][t], int(z["before_active_agents"][t]))
        got = env.get_state()
        for k in INT_STATE:
            assert int(got[k]) == int(z[f"after_{k}"][t]), (name, t, k)
        assert mask == int(z["after_active_agents"][t]), (name, t)
        assert np.array_equal(term, z["terminated"][t].astype(bool)) and np.array_equal(trunc, z["truncated"][t].astype(bool)), (name, t)
        assert bool(z["term_all"][t]) == (mask == 0) and bool(z["trunc_all"][t]) == (mask == 0)
        for k in FLOAT_STATE:
            ref = z[f"after_{k}"][t]
            scale = max(1.0, float(np.nanmax(np.abs(np.where(np.isfinite(ref), ref, 0.0)))))
            assert max_err(got[k], ref) <= TOL * scale, (name, t, k)
        assert max_err(obs, z["obs"][t]) <= 2e-7 * max(1.0, float(np.max(np.abs(z["obs"][t])))), (name, t, "obs")
        assert max_err(rew, z["reward"][t]) <= TOL * max(1.0, float(np.max(np.abs(z["reward"][t])))), (name, t, "reward")
        masks.add(int(z["before_active_agents"][t]))
    assert len(masks) >= 1
