"""GPU parity of the MARL packing (rllib_envs semantics) against the golden rollouts recorded from
`MARLCattleAviary` and against the CPU oracle's `step_marl`.

Same teacher-forced protocol as tests/test_cuda_parity.py: recorded step t goes into environment t, one launch
steps them all.  Per-drone rewards and flags are compared slot by slot; column D of the flag arrays is RLlib's
"__all__" entry (rllib_envs/BaseAviary.py:428,430).
"""
import numpy as np
import pytest

from _golden import FLOAT_STATE, INT_STATE, Golden, golden_names, max_err
from test_cuda_parity import FP32_FLOOR, FP32_REL, _mask_near_ties

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import cattle_oracle as co  # noqa: E402

D = 12


def _env_for(g, precision, n_envs=None, **kw):
    import rl_cattle_herding_b200 as pkg
    m = g.meta
    args = dict(num_drones=D, num_cattle=m["n_cattle"], ctrl_freq=m["ctrl_freq"], pyb_freq=m["pyb_freq"], act=m["act"],
                start_level=m["start_level"], auto_reset=False, precision=precision, spawn_table=g.table, device=0,
                packing="marl")
    args.update(kw)
    return pkg.CattleHerdingBatchedEnv(g.steps if n_envs is None else n_envs, **args)


def _check_rewards(rew, ref, tol, tag):
    err = np.abs(rew.astype(np.float64) - ref)
    bound = tol * (1.0 + np.abs(ref))
    assert np.all(err <= bound), (tag, float(err.max()), np.argwhere(err > bound)[:4])


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
@pytest.mark.parametrize("name", golden_names("marl"))
def test_marl_teacher_forced(name, precision):
    g = Golden(name)
    env = _env_for(g, precision)
    assert env.obs.shape == (g.steps, D, g.meta["obs_len"]) and env.reward.shape == (g.steps, D)
    assert env.terminated.shape == (g.steps, D + 1)
    env.set_state({k: g.z[f"before_{k}"] for k in FLOAT_STATE + INT_STATE})
    obs, rew, term, trunc = env.step(torch.from_numpy(np.ascontiguousarray(g.z["action"])).cuda())
    torch.cuda.synchronize()
    got = env.get_state()
    obs, rew, term, trunc = obs.cpu().numpy(), rew.cpu().numpy(), term.cpu().numpy(), trunc.cpu().numpy()
    for k in INT_STATE:
        assert np.array_equal(got[k].astype(np.int64), g.z[f"after_{k}"].astype(np.int64)), (name, k)
    assert np.array_equal(term[:, :D], g.z["terminated"]), (name, "terminated")
    assert np.array_equal(trunc[:, :D], g.z["truncated"]), (name, "truncated")
    assert np.array_equal(term[:, D].astype(bool), g.z["term_all"]) and np.array_equal(trunc[:, D].astype(bool), g.z["trunc_all"])
    for k in FLOAT_STATE:
        ref = g.z[f"after_{k}"]
        scale = max(1.0, float(np.max(np.abs(np.where(np.isfinite(ref), ref, 0.0)))))
        tol = 1e-9 * scale if precision == "fp64" else FP32_REL * scale + FP32_FLOOR.get(k, 1e-6)
        assert max_err(got[k], ref) <= tol, (name, precision, k, max_err(got[k], ref), tol)
    ref_obs = g.z["obs"]
    scale = max(1.0, float(np.max(np.abs(ref_obs))))
    if precision == "fp32":
        obs = _mask_near_ties(g, obs, ref_obs)
    assert max_err(obs, ref_obs) <= (2e-7 * scale if precision == "fp64" else FP32_REL * scale + 1e-6), (name, precision, "obs")
    # float32 reward tensor: 2^-24 relative in the fp64 build; fp32 build as in test_cuda_parity (approach gain)
    _check_rewards(rew, g.z["reward"], 1.2e-7 if precision == "fp64" else 5e-5, (name, precision))
    env.close()


@pytest.mark.parametrize("level", [0, 2, 4, 5, 6])
def test_marl_random_states_vs_oracle(level):
    """States the fixtures do not reach: 64 envs per level, tight formations around the herd so that success,
    mid-step level-ups and every truncation cause fire; fp64 build against the oracle's step_marl."""
    import rl_cattle_herding_b200 as pkg
    E, M, N = 64, 16, 4
    table = pkg.spawn.generate(5, M, seed=3)
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, ctrl_freq=30, start_level=level, fixed_num_drones=N,
                                      auto_reset=False, precision="fp64", spawn_table=table, device=0, packing="marl", seed=4)
    env.reset()
    st = env.get_state()
    rng = np.random.default_rng(level)
    cent = st["cattle_pos"][:, :, :2].mean(axis=1)
    for e in range(E):
        r = rng.choice([0.5, 1.0, 3.0, 6.0, 9.0])
        ang = rng.uniform(0, 2 * np.pi) + np.arange(N) * 2 * np.pi / N
        off = rng.choice([0.0, 0.0, 4.0, 16.0])
        st["pos"][e, :, 0] = cent[e, 0] + off + r * np.cos(ang) + rng.normal(0, 0.05, N)
        st["pos"][e, :, 1] = cent[e, 1] + r * np.sin(ang) + rng.normal(0, 0.05, N)
        st["pos"][e, :, 2] = 2.0 + rng.choice([0.0, 0.0, 0.0, 1.5]) * (np.arange(N) == rng.integers(N))
        st["tally"][e] = rng.choice([0, 1, 1000])          # 1000: the next success levels the env up mid-loop
        st["spacing_clock"][e] = rng.choice([0.0, 10.0])
        st["prev_cent_dist"][e] = rng.choice([np.nan, 3.0])
        st["step_counter"][e] = rng.choice([0, 10, 30 * 80 + 5])
    env.set_state(st)
    st = env.get_state()
    act = rng.uniform(-1, 1, (E, N, 4)).astype(np.float32)
    obs, rew, term, trunc = [x.cpu().numpy() for x in env.step(torch.from_numpy(act).cuda())]
    after = env.get_state()
    seen_term = seen_trunc = seen_up = 0
    for e in range(E):
        o = co.OracleEnv(M, ctrl_freq=30, start_level=level, spawn_table=table)
        full = {}
        for k in co.STATE_FIELDS:
            a = np.asarray(st[k][e])
            if a.ndim == 2 and a.shape[0] == N and k not in ("cattle_pos", "cattle_vel"):
                pad = np.zeros((12, a.shape[1])); pad[:N] = a; a = pad
            full[k] = a
        o.set_state(full)
        oobs, orew, oterm, otrunc = o.step_marl(act[e])
        assert np.array_equal(term[e, :N].astype(bool), oterm[:N]) and np.array_equal(trunc[e, :N].astype(bool), otrunc[:N]), e
        assert bool(term[e, N]) == bool(oterm[:N].all()) and bool(trunc[e, N]) == bool(otrunc[:N].all()), e
        assert int(after["level"][e]) == o.level and int(after["tally"][e]) == o.tally, e
        assert int(after["step_counter"][e]) == o.step_counter
        assert max_err(after["spacing_clock"][e], o.spacing_clock) <= 1e-12
        assert max_err(obs[e, :N], oobs[:N]) <= 4e-6
        _check_rewards(rew[e, :N], orew[:N], 2e-7, (level, e))
        seen_term += int(oterm[:N].any()); seen_trunc += int(otrunc[:N].any()); seen_up += int(o.level != level)
    assert seen_trunc > 0 and (seen_term > 0 or level == 5)
    env.close()


def test_marl_auto_reset_and_host_path():
    """auto_reset fires on either "__all__" flag; the host-buffer call returns the same arrays as the device call."""
    import rl_cattle_herding_b200 as pkg
    E, M, N = 128, 16, 3
    table = pkg.spawn.generate(9, M, seed=1)
    mk = lambda: pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, ctrl_freq=30, start_level=0, fixed_num_drones=N,
                                             auto_reset=True, spawn_table=table, device=0, packing="marl", seed=2)
    a, b = mk(), mk()
    a.reset(); b.reset_host()
    rng = np.random.default_rng(0)
    n_done = 0
    for t in range(40):
        act = rng.uniform(-1, 1, (E, N, 4)).astype(np.float32)
        epi_before = a.get_state()["episode"].copy()
        oa, ra, ta, ua = [x.cpu().numpy() for x in a.step(torch.from_numpy(act).cuda())]
        ob, rb, tb, ub, idx, tobs = b.step_host(act)
        assert np.array_equal(oa, ob) and np.array_equal(ra, rb) and np.array_equal(ta, tb) and np.array_equal(ua, ub)
        done = (ta[:, N] | ua[:, N]).astype(bool)
        assert np.array_equal(a.get_state()["episode"] - epi_before, done.astype(np.int32)), t
        assert sorted(idx.tolist()) == np.nonzero(done)[0].tolist()
        n_done += int(done.sum())
    assert n_done > 20
    a.close(); b.close()


# ---- the reference's RLlib wrapper, bug for bug (ChsParams.marl_wrapper_quirks) ---------------------------------------
WRAPPER_GOLDEN = ["w_3d20c_uniform", "w_l0_hold_levelup", "w_l2_approach", "w_l4_surround"]


class _WGolden(Golden):
    def __init__(self, name):
        import json, os
        from _golden import GOLDEN_DIR
        z = np.load(os.path.join(GOLDEN_DIR, "wrapper", name + ".npz"))
        self.name, self.z = name, {k: z[k] for k in z.files}
        self.meta = json.loads(str(self.z["meta_json"]))
        self.steps, self.table = self.z["action"].shape[0], self.z["spawn_table"]


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
@pytest.mark.parametrize("name", WRAPPER_GOLDEN)
def test_rllib_wrapper_quirks_teacher_forced(name, precision):
    """What RLlib sees through the REFERENCE wrapper (rllib_envs/marl_wrapper.py:77-119): rollouts recorded through that
    wrapper (oracle/ref_harness/make_wrapper_golden.py), step t injected into env t together with the wrapper's agent
    list, one launch with `marl_wrapper_quirks`: rewards / flags of the second invocation, the state it leaves (spacing
    clock, tally, level, prev_cent_dist) and the agent list after the step."""
    g = _WGolden(name)
    env = _env_for(g, precision, marl_wrapper_quirks=True)
    ints = INT_STATE + ("active_agents",)
    env.set_state({k: g.z[f"before_{k}"] for k in FLOAT_STATE + ints})
    obs, rew, term, trunc = env.step(torch.from_numpy(np.ascontiguousarray(g.z["action"])).cuda())
    torch.cuda.synchronize()
    got = env.get_state()
    obs, rew, term, trunc = obs.cpu().numpy(), rew.cpu().numpy(), term.cpu().numpy(), trunc.cpu().numpy()
    for k in ints:
        assert np.array_equal(got[k].astype(np.int64), g.z[f"after_{k}"].astype(np.int64)), (name, k, got[k], g.z[f"after_{k}"])
    assert np.array_equal(term[:, :D], g.z["terminated"]), (name, "terminated")
    assert np.array_equal(trunc[:, :D], g.z["truncated"]), (name, "truncated")
    assert np.array_equal(term[:, D].astype(bool), g.z["term_all"].astype(bool)) and np.array_equal(trunc[:, D].astype(bool), g.z["trunc_all"].astype(bool))
    for k in FLOAT_STATE:
        ref = g.z[f"after_{k}"]
        scale = max(1.0, float(np.max(np.abs(np.where(np.isfinite(ref), ref, 0.0)))))
        tol = 1e-9 * scale if precision == "fp64" else FP32_REL * scale + FP32_FLOOR.get(k, 1e-6)
        assert max_err(got[k], ref) <= tol, (name, precision, k, max_err(got[k], ref), tol)
    # observations: the wrapper returns rows of the agents in its list only (zeros elsewhere in the recording)
    listed = ((g.z["before_active_agents"][:, None] >> np.arange(D)[None, :]) & 1).astype(bool)
    ref_obs = g.z["obs"]
    scale = max(1.0, float(np.max(np.abs(ref_obs))))
    if precision == "fp32":
        obs = _mask_near_ties(g, obs, np.where(listed[:, :, None], ref_obs, obs))
    assert max_err(np.where(listed[:, :, None], obs, 0.0), ref_obs) <= (2e-7 * scale if precision == "fp64" else FP32_REL * scale + 1e-6)
    _check_rewards(rew, g.z["reward"], 1.2e-7 if precision == "fp64" else 5e-5, (name, precision))
    env.close()


def test_rllib_wrapper_quirks_free_running():
    """`RLlibMultiAgentWrapper` with `reference_wrapper_quirks`: replays the recorded run from its first state through the
    dict API (agent list kept by the wrapper AND by the kernel) and lands on the reference wrapper's outputs."""
    import rl_cattle_herding_b200 as pkg
    g = _WGolden("w_l0_hold_levelup")
    m = g.meta
    w = pkg.RLlibMultiAgentWrapper(dict(num_drones=D, num_cattle=m["n_cattle"], ctrl_freq=m["ctrl_freq"], fixed_num_drones=m["n_drones"],
                                        curriculum_level=m["start_level"], precision="fp64", spawn_table=g.table,
                                        reference_wrapper_quirks=True))
    w.reset()
    w.env.sim.set_state({k: g.z[f"before_{k}"][:1] for k in FLOAT_STATE + INT_STATE})
    for t in range(g.steps):
        assert sum(1 << int(a.split("_")[1]) for a in w.agents) == int(g.z["before_active_agents"][t]), t
        obs, rew, term, trunc, _ = w.step({aid: g.z["action"][t][int(aid.split("_")[1])] for aid in w.agents})
        for aid in rew:
            i = int(aid.split("_")[1])
            assert abs(rew[aid] - g.z["reward"][t, i]) <= 2e-6 * (1 + abs(g.z["reward"][t, i])), (t, aid)
            assert term[aid] == bool(g.z["terminated"][t, i]) and trunc[aid] == bool(g.z["truncated"][t, i]), (t, aid)
        assert term["__all__"] == bool(g.z["term_all"][t]) and trunc["__all__"] == bool(g.z["trunc_all"][t])
        assert not g.z["reset_after"][t]
    w.close()
