"""GPU: edge shapes of the batched step against the float64 oracle -- ragged batches (E not a multiple of the envs per
CTA), a single env, one cow, herds larger than one 32-cow chunk, the largest fleet (12 drone slots) with the number of
drones drawn per episode, both step modes.  A few steps from the device's own reset state, every env checked."""
import numpy as np
import pytest

from _golden import max_err  # noqa: F401  (puts oracle/ on sys.path)

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import cattle_oracle as co  # noqa: E402


DRONE_FIELDS = ("pos", "quat", "rpy", "vel", "ang_v", "rpy_rates", "last_rpm", "pid_ipos", "pid_last_rpy", "pid_irpy")


def _pad12(key, a):
    """Oracle state arrays have 12 drone slots; the device env has D."""
    a = np.asarray(a)
    if key in DRONE_FIELDS and a.shape[0] < 12:
        out = np.zeros((12, a.shape[1]))
        out[: a.shape[0]] = a
        return out
    return a


CASES = [
    # E, D, M, envs/CTA, threads, fixed_n, level
    (37, 3, 20, 8, 128, 3, 7),      # ragged: last CTA holds 5 of 8 envs
    (1, 3, 20, 4, 64, 3, 7),        # a single env
    (5, 3, 1, 2, 64, 3, 7),         # one cow: no alpha pairs, 15 empty cattle slots in the observation
    (9, 4, 33, 4, 128, 4, 5),       # herd one cow beyond a 32-cow mask chunk; level with spacing + effectiveness success
    (6, 12, 16, 2, 256, 0, 7),      # 12 slots, n drawn per episode in the level's [4, 12]
    (3, 8, 100, 1, 256, 8, 6),      # BASELINE config C shape, one env per CTA (full build of the kernel)
]


@pytest.mark.parametrize("mode", ["fused", "split"])
@pytest.mark.parametrize("case", CASES, ids=lambda c: "E%d_D%d_M%d_G%d" % c[:4])
def test_edge_shapes_against_oracle(case, mode):
    import rl_cattle_herding_b200 as pkg
    E, D, M, G, TH, fixed_n, level = case
    table = pkg.spawn.generate(5, M, seed=2)
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=D, num_cattle=M, ctrl_freq=30, fixed_num_drones=fixed_n, start_level=level,
                                      spawn_table=table, device=0, seed=5, precision="fp64", auto_reset=False,
                                      envs_per_block=G, threads_per_block=TH, step_mode=mode)
    env.reset()
    rng = np.random.default_rng(E * 100 + D)
    worst = 0.0
    for step in range(3):
        st = env.get_state()
        act = rng.uniform(-1, 1, (E, D, 4)).astype(np.float32)
        obs, rew, term, trunc = env.step(torch.from_numpy(act).cuda())
        torch.cuda.synchronize()
        obs, rew, term, trunc = obs.cpu().numpy(), rew.cpu().numpy(), term.cpu().numpy(), trunc.cpu().numpy()
        for e in range(E):
            n = int(st["n_drones"][e])
            assert (fixed_n and n == fixed_n) or (not fixed_n and 4 <= n <= 12)
            o = co.OracleEnv(M, ctrl_freq=30, spawn_table=table, start_level=level)
            o.set_state({k: _pad12(k, st[k][e]) for k in co.STATE_FIELDS})
            oobs, orew, oterm, otrunc, _ = o.step(act[e, :n])
            worst = max(worst, float(np.max(np.abs(oobs - obs[e]))), abs(orew - float(rew[e])))
            assert bool(term[e]) == oterm and bool(trunc[e]) == otrunc, (step, e)
    assert worst < 1e-6, worst          # float64 build; the obs / reward tensors are float32
    env.close()


@pytest.mark.parametrize("shape", [(3, 20), (8, 100), (4, 33)], ids=lambda s: "D%d_M%d" % s)
def test_alpha_range_table_equals_computed_ranges(shape, monkeypatch):
    """The per-spawn-row tables of the alpha term (built by chs_set_spawn_table) against the distance pass they replace, over
    free-running steps with resets in between.  The in-range MASKS alone give the same herd velocities bit for bit.  With the
    WEIGHT table the gradient sum of a cow is precomputed by another kernel (same expressions, another compilation: fp32
    contraction may differ in the last bit), so velocities agree to rounding; everything a step returns is still bit-equal
    (the herd does not move on the DYN path, so velocities feed nothing back)."""
    import rl_cattle_herding_b200 as pkg
    D, M = shape
    E = 64
    table = pkg.spawn.generate(6, M, seed=4)

    def make():
        return pkg.CattleHerdingBatchedEnv(E, num_drones=D, num_cattle=M, ctrl_freq=30, fixed_num_drones=D, spawn_table=table,
                                           device=0, seed=9, auto_reset=True)
    with_weights = make()
    monkeypatch.setenv("CHS_NO_ALPHA_WEIGHTS", "1")     # both read by chs_set_spawn_table
    with_table = make()
    monkeypatch.setenv("CHS_NO_ALPHA_TABLE", "1")
    computed = make()
    monkeypatch.delenv("CHS_NO_ALPHA_TABLE")
    monkeypatch.delenv("CHS_NO_ALPHA_WEIGHTS")
    for env in (with_weights, with_table, computed):
        env.reset()
    gen = torch.Generator("cuda").manual_seed(2)
    for t in range(40):
        act = torch.rand((E, D, 4), device="cuda", generator=gen) * 2 - 1
        a = [x.clone() for x in with_table.step(act)]
        w = [x.clone() for x in with_weights.step(act)]
        b = computed.step(act)
        assert all(torch.equal(x, y) for x, y in zip(a, b)), t
        assert all(torch.equal(x, y) for x, y in zip(w, b)), t
    sa, sw, sb = with_table.get_state(), with_weights.get_state(), computed.get_state()
    assert np.array_equal(sa["cattle_vel"], sb["cattle_vel"]) and np.array_equal(sa["cattle_pos"], sb["cattle_pos"])
    assert np.array_equal(sw["cattle_pos"], sb["cattle_pos"])
    assert np.allclose(sw["cattle_vel"], sb["cattle_vel"], rtol=0, atol=2e-6), np.abs(sw["cattle_vel"] - sb["cattle_vel"]).max()
    assert with_table.stats()[0] > 0          # episodes ended and restarted on other spawn rows along the way
    with_weights.close(); with_table.close(); computed.close()
