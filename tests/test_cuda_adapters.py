"""GPU: the SB3-style VecEnv adapter and the single-env facades with the reference's class names."""
import numpy as np
import pytest

from _golden import max_err

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import cattle_oracle as co  # noqa: E402


def test_vec_env_contract():
    import rl_cattle_herding_b200 as pkg
    E = 64
    venv = pkg.GpuCattleVecEnv(E, num_drones=3, num_cattle=20, ctrl_freq=30, fixed_num_drones=3, device=0, seed=2,
                               autotune_launch=True)      # the tuned launch shape must not change a single bit
    twin = pkg.CattleHerdingBatchedEnv(E, num_drones=3, num_cattle=20, ctrl_freq=30, fixed_num_drones=3, device=0, seed=2,
                                       spawn_table=pkg.spawn.generate(100, 20, seed=0))
    assert venv.num_envs == E and venv.observation_space.shape == (12, 66) and venv.action_space.shape == (3, 4)
    assert venv.observation_space.dtype == np.float32 and float(venv.action_space.low.min()) == -1.0
    obs = venv.reset()
    tobs = twin.reset().cpu().numpy()
    assert obs.shape == (E, 12, 66) and obs.dtype == np.float32 and np.array_equal(obs, tobs)
    rng = np.random.default_rng(4)
    finished = 0
    for t in range(40):
        act = rng.uniform(-1, 1, (E, 3, 4)).astype(np.float32)
        obs, rew, dones, infos = venv.step(act)
        o2, r2, te2, tr2 = twin.step(torch.from_numpy(act).cuda())
        assert np.array_equal(obs, o2.cpu().numpy()) and np.array_equal(rew, r2.cpu().numpy())
        assert dones.dtype == bool and np.array_equal(dones, (te2 | tr2).bool().cpu().numpy())
        assert len(infos) == E
        for e in range(E):
            if dones[e]:
                finished += 1
                assert np.array_equal(infos[e]["terminal_observation"], twin.terminal_obs[e].cpu().numpy())
                assert infos[e]["TimeLimit.truncated"] == (bool(tr2[e]) and not bool(te2[e]))
            else:
                assert "terminal_observation" not in infos[e]
    assert finished > 20
    st = venv.episode_stats()
    assert st["episodes"] == finished and st["env_steps"] == 40 * E
    assert venv.get_attr("CTRL_FREQ") == [30] * E and venv.env_is_wrapped(object) == [False] * E
    venv.close(); twin.close()


def test_vec_env_rejects_what_the_platform_cannot_do():
    import rl_cattle_herding_b200 as pkg
    with pytest.raises(ValueError, match="DYN"):
        pkg.GpuCattleVecEnv(4, physics=pkg.Physics.PYB)
    with pytest.raises(ValueError, match="gui"):
        pkg.GpuCattleVecEnv(4, gui=True)


@pytest.mark.parametrize("precision,tol", [("fp64", 1e-9), ("fp32", 5e-5)])
def test_cattle_aviary_facade_against_oracle(precision, tol):
    """Reference-shaped single env: CattleAviary(...).reset()/step() vs the oracle stepping from the same state."""
    import rl_cattle_herding_b200 as pkg
    table = pkg.spawn.generate(5, 20, seed=0)
    env = pkg.CattleAviary(num_drones=3, num_cattle=20, physics=pkg.Physics.DYN, pyb_freq=240, ctrl_freq=30,
                           fixed_num_drones=3, spawn_table=table, precision=precision, device=0, seed=9)
    obs, info = env.reset()
    assert obs.shape == (12, 66) and obs.dtype == np.float32 and info == {"answer": 42}
    assert env.NUM_DRONES == 3 and env.EPISODE_LEN_SEC == 80 and env.CTRL_TIMESTEP == 1 / 30
    assert env.action_space.shape == (3, 4) and env.observation_space.shape == (12, 66)
    rng = np.random.default_rng(1)
    for t in range(12):
        st = env.sim.get_state()
        o = co.OracleEnv(20, ctrl_freq=30, spawn_table=table)
        o.set_state({k: _pad(st[k][0]) for k in co.STATE_FIELDS})
        a = rng.uniform(-1, 1, (3, 4)).astype(np.float32)
        obs, rew, term, trunc, info = env.step(a)
        oobs, orew, oterm, otrunc, _ = o.step(a)
        assert isinstance(rew, float) and isinstance(term, bool) and isinstance(trunc, bool)
        assert (term, trunc) == (oterm, otrunc)
        assert max_err(obs, oobs) <= max(tol, 3e-7) * 10 and abs(rew - orew) <= max(tol, 1.2e-7)
        if term or trunc:
            env.reset()
    sv = env._getDroneStateVector(0)
    assert sv.shape == (20,) and env._getCowStateVector(3).shape == (16,)
    assert env.HerdCentroid().shape == (3,) and env.DroneCentroid()[2] == 0.95
    env.close()


def _pad(a):
    a = np.asarray(a)
    if a.ndim == 2 and a.shape[0] < 12 and a.shape[1] in (3, 4):
        out = np.zeros((12, a.shape[1]))
        out[: a.shape[0]] = a
        return out
    return a


def test_marl_facade_and_rllib_wrapper_against_oracle():
    """MARLCattleAviary dict contract and the agent_<i> wrapper (rllib_envs/marl_wrapper.py) vs the oracle's step_marl."""
    import rl_cattle_herding_b200 as pkg
    M, N = 16, 3
    table = pkg.spawn.generate(5, M, seed=0)
    cfg = dict(num_drones=N, num_cattle=M, ctrl_freq=30, fixed_num_drones=N, spawn_table=table, device=0, seed=1, precision="fp64")
    env = pkg.MARLCattleAviary(**cfg)
    wrap = pkg.RLlibMultiAgentWrapper(cfg)
    assert env.observation_space.shape == (66,) and env.action_space.shape == (4,)
    obs, info = env.reset()
    wobs, winfo = wrap.reset()
    assert sorted(obs) == [0, 1, 2] and info == {"__all__": {}} and wrap.agents == ["agent_0", "agent_1", "agent_2"]
    assert all(np.array_equal(obs[i], wobs[f"agent_{i}"]) for i in range(N))
    o = co.OracleEnv(M, ctrl_freq=30, start_level=0, spawn_table=table)
    st = env.sim.get_state()
    full = {}
    for k in co.STATE_FIELDS:
        a = np.asarray(st[k][0])
        if a.ndim == 2 and a.shape[0] == N and k not in ("cattle_pos", "cattle_vel"):
            pad = np.zeros((12, a.shape[1])); pad[:N] = a; a = pad
        full[k] = a
    o.set_state(full)
    rng = np.random.default_rng(0)
    for t in range(12):
        act = rng.uniform(-1, 1, (N, 4)).astype(np.float32)
        obs, rew, done, trunc, info = env.step(act)
        wo, wr, wd, wt, wi = wrap.step({f"agent_{i}": act[i] for i in range(N)})
        oobs, orew, oterm, otrunc = o.step_marl(act)
        assert set(done) == {0, 1, 2, "__all__"} and set(trunc) == {0, 1, 2, "__all__"}
        for i in range(N):
            assert max_err(obs[i], oobs[i]) <= 4e-6 and abs(rew[i] - orew[i]) <= 2e-7 * (1 + abs(orew[i]))
            assert done[i] == bool(oterm[i]) and trunc[i] == bool(otrunc[i])
            assert np.array_equal(wo[f"agent_{i}"], obs[i]) and wr[f"agent_{i}"] == rew[i]
        assert done["__all__"] == bool(oterm[:N].all()) and trunc["__all__"] == bool(otrunc[:N].all())
        assert wd["__all__"] == (len(wrap.agents) == 0)
        if any(trunc.values()) or any(done.values()):
            break
    env.close(); wrap.close()
