"""GPU: the opt-in non-reference extensions against the CPU restatement that defines
them (oracle/cattle_oracle.py): aerodynamic extras inside the DYN step and the kinematic herd."""
import numpy as np
import pytest

from _golden import max_err

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import cattle_oracle as co  # noqa: E402


def _oracle_from(st, e, N, M, table, **kw):
    o = co.OracleEnv(M, ctrl_freq=30, spawn_table=table, **kw)
    full = {}
    for k in co.STATE_FIELDS:
        a = np.asarray(st[k][e])
        if a.ndim == 2 and a.shape[0] == N and k not in ("cattle_pos", "cattle_vel"):
            pad = np.zeros((12, a.shape[1])); pad[:N] = a; a = pad
        full[k] = a
    o.set_state(full)
    return o


@pytest.mark.parametrize("precision,tol", [("fp64", 1e-10), ("fp32", 2e-5)])
@pytest.mark.parametrize("flags", [1, 2, 4, 7])
def test_aero_extras_match_the_cpu_restatement(flags, precision, tol):
    import rl_cattle_herding_b200 as pkg
    E, M, N = 32, 16, 4
    table = pkg.spawn.generate(5, M, seed=2)
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, ctrl_freq=30, fixed_num_drones=N, auto_reset=False,
                                      precision=precision, spawn_table=table, device=0, seed=3, aero=flags)
    env.reset()
    rng = np.random.default_rng(flags)
    st = env.get_state()
    # stacked / low / fast drones so that every model has something to act on
    st["pos"][:, :, 2] = rng.choice([0.06, 0.2, 0.45, 0.9], size=(E, N))
    st["pos"][:, 1, :2] = st["pos"][:, 0, :2] + rng.normal(0, 0.05, (E, 2))     # drone 1 hovers 0.4-0.6 m above drone 0
    st["pos"][:, 0, 2] = rng.choice([0.06, 0.2], size=E)
    st["pos"][:, 1, 2] = st["pos"][:, 0, 2] + rng.uniform(0.4, 0.6, E)          # (the model is singular as dz -> 0)
    st["vel"][:] = rng.normal(0, 1.0, (E, N, 3))
    st["last_rpm"][:] = rng.uniform(8000, 20000, (E, N, 4))
    env.set_state(st)
    st = env.get_state()
    oracles = [_oracle_from(st, e, N, M, table, aero_flags=flags) for e in range(E)]
    plain = [_oracle_from(st, e, N, M, table) for e in range(E)]
    moved = 0.0
    for t in range(3):
        act = rng.uniform(-1, 1, (E, N, 4)).astype(np.float32)
        obs, rew, term, trunc = [x.cpu().numpy() for x in env.step(torch.from_numpy(act).cuda())]
        got = env.get_state()
        for e in range(E):
            oobs, orew, oterm, otrunc, _ = oracles[e].step(act[e])
            plain[e].step(act[e])
            for k, ref in (("pos", oracles[e].pos), ("vel", oracles[e].vel), ("quat", oracles[e].quat), ("rpy_rates", oracles[e].rpy_rates)):
                scale = max(1.0, float(np.abs(ref[:N]).max()))
                assert max_err(got[k][e], ref[:N]) <= tol * scale * (3 if k == "rpy_rates" else 1), (flags, precision, t, e, k)
            assert bool(term[e]) == oterm and bool(trunc[e]) == otrunc
            moved = max(moved, float(np.abs(oracles[e].vel[:N] - plain[e].vel[:N]).max()))
    assert moved > 1e-3          # the extras change the trajectory (otherwise the test shows nothing)
    env.close()


@pytest.mark.parametrize("packing", ["ctde", "marl"])
@pytest.mark.parametrize("precision,tol", [("fp64", 1e-10), ("fp32", 2e-5)])
def test_cattle_integrate_matches_the_cpu_restatement(precision, tol, packing):
    import rl_cattle_herding_b200 as pkg
    E, M, N = 24, 20, 3
    table = pkg.spawn.generate(5, M, seed=2)
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, ctrl_freq=30, fixed_num_drones=N, auto_reset=False,
                                      precision=precision, spawn_table=table, device=0, seed=3, cattle_integrate=True,
                                      packing=packing, start_level=4)
    env.reset()
    st = env.get_state()
    oracles = [_oracle_from(st, e, N, M, table, cattle_integrate=True, start_level=4) for e in range(E)]
    start = st["cattle_pos"].copy()
    rng = np.random.default_rng(0)
    for t in range(6):
        act = rng.uniform(-1, 1, (E, N, 4)).astype(np.float32)
        out = [x.cpu().numpy() for x in env.step(torch.from_numpy(act).cuda())]
        got = env.get_state()
        for e in range(E):
            o = oracles[e]
            res = o.step(act[e]) if packing == "ctde" else o.step_marl(act[e])
            assert max_err(got["cattle_pos"][e][:, :2], o.cattle_pos[:, :2]) <= tol * 20, (t, e)
            assert max_err(got["cattle_vel"][e][:, :2], o.cattle_vel[:, :2]) <= max(tol, 4e-6), (t, e)
            oobs = res[0]
            rows = 12 if packing == "ctde" else N
            got_obs, ref_obs = out[0][e][:rows].copy(), oobs[:rows].copy()
            if precision == "fp32":      # the line spawn leaves the middle drone equidistant from its neighbours:
                got_obs[:, 10:14] = 0    # sub-fp32 ranking ties are covered by tests/test_cuda_parity.py
                ref_obs[:, 10:14] = 0
            assert max_err(got_obs, ref_obs) <= max(tol * 20, 4e-6), (t, e)
            orew = np.atleast_1d(res[1])[: (1 if packing == "ctde" else N)]
            rew = np.atleast_1d(out[1][e])[: len(orew)]
            assert np.all(np.abs(rew - orew) <= (2e-7 if precision == "fp64" else 5e-5) * (1 + np.abs(orew))), (t, e)
    assert np.abs(env.get_state()["cattle_pos"] - start).max() > 1e-2       # the herd moved
    env.close()
