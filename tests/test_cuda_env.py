"""GPU: reset, auto-reset, host-buffer path, launch-shape and sharding invariance of the batched simulator."""
import numpy as np
import pytest

from _golden import max_err

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import cattle_oracle as co  # noqa: E402  (tests/_golden.py put oracle/ on sys.path)


def _pkg():
    import rl_cattle_herding_b200 as pkg
    return pkg


def _make(E, **kw):
    pkg = _pkg()
    M = kw.pop("num_cattle", 20)
    table = kw.pop("spawn_table", None)
    if table is None:
        table = pkg.spawn.generate(7, M, seed=0)
    args = dict(num_drones=3, num_cattle=M, fixed_num_drones=3, spawn_table=table, device=0, seed=11)
    args.update(kw)
    return pkg.CattleHerdingBatchedEnv(E, **args), table


def _pad12(a):
    a = np.asarray(a)
    if a.ndim == 2 and a.shape[0] < 12 and a.shape[1] in (3, 4):
        out = np.zeros((12, a.shape[1]))
        out[: a.shape[0]] = a
        return out
    return a


@pytest.mark.parametrize("precision,tol", [("fp64", 1e-12), ("fp32", 2e-6)])
@pytest.mark.parametrize("variable_n", [False, True])
def test_reset_matches_oracle(precision, tol, variable_n):
    """Device reset (Philox draws) == oracle reset fed with the same draws (BaseAviary.py:280-331,547-637)."""
    from rl_cattle_herding_b200 import philox as ph
    E, M, seed, off = 48, 16 if variable_n else 20, 5, 100
    kw = dict(num_cattle=M, precision=precision, seed=seed, env_id_offset=off)
    if variable_n:
        kw.update(num_drones=12, fixed_num_drones=0, start_level=7)      # 4..12 drones, two-row spawn
    env, table = _make(E, **kw)
    for episode in (1, 2):
        obs = env.reset().cpu().numpy()
        st = env.get_state()
        for e in range(E):
            n = ph.num_drones(seed, off + e, episode, 4, 12) if variable_n else 3
            o = co.OracleEnv(M, spawn_table=table)
            o.spawn_idx = (off + e + episode - 1) % table.shape[0]
            oobs = o.reset(n, ph.cow_uniforms(seed, off + e, episode, M))
            assert int(st["n_drones"][e]) == n and int(st["spawn_idx"][e]) == o.spawn_idx
            assert int(st["step_counter"][e]) == 0 and int(st["episode"][e]) == episode
            D = env.max_drones
            assert max_err(st["pos"][e], o.pos[:D]) <= tol * 10 and max_err(st["quat"][e], o.quat[:D]) == 0
            assert max_err(st["cattle_pos"][e], o.cattle_pos) <= tol * 20
            assert max_err(st["cattle_vel"][e], o.cattle_vel) <= tol
            assert max_err(obs[e], oobs) <= max(tol * 20, 2e-6), (e, max_err(obs[e], oobs))
    env.close()


def test_auto_reset_equals_manual_reset():
    """In-kernel reset of finished envs == masked chs_reset on a twin without auto-reset, bit for bit; the
    terminal observation is the twin's last observation (SB3 VecEnv contract)."""
    E = 256
    a, _ = _make(E, auto_reset=True)
    b, _ = _make(E, auto_reset=False)
    a.reset(); b.reset()
    gen = torch.Generator("cuda").manual_seed(3)
    n_done = 0
    for t in range(40):
        act = torch.rand((E, 3, 4), device="cuda", generator=gen) * 2 - 1
        oa, ra, ta, ua = [x.clone() for x in a.step(act)]
        ob, rb, tb, ub = [x.clone() for x in b.step(act)]
        assert torch.equal(ra, rb) and torch.equal(ta, tb) and torch.equal(ua, ub), t
        done = (tb | ub).bool()
        if done.any():
            assert torch.equal(a.terminal_obs[done], ob[done]), t
            ob = b.reset(mask=done).clone()
            n_done += int(done.sum())
        assert torch.equal(oa, ob), t
        sa, sb = a.get_state(), b.get_state()
        for k in sa:
            assert np.array_equal(sa[k], sb[k], equal_nan=True), (t, k)
    assert n_done > 50
    st = a.stats()
    assert st[0] == n_done and st[10] == 40 * E
    assert st[3:9].sum() == n_done      # every finished episode has exactly one cause
    a.close(); b.close()


def test_state_survives_reset_like_the_reference():
    """PID integrators, prev_cent_dist, spacing clock and curriculum survive a reset."""
    env, _ = _make(8, auto_reset=False, start_level=0)
    env.reset()
    st = env.get_state()
    st["pid_ipos"][:] = 0.125; st["pid_last_rpy"][:] = -0.25; st["pid_irpy"][:] = 0.5
    st["prev_cent_dist"][:] = 4.5; st["spacing_clock"][:] = 1.25; st["tally"][:] = 7; st["level"][:] = 2
    env.set_state(st)
    env.reset()
    s2 = env.get_state()
    for k in ("pid_ipos", "pid_last_rpy", "pid_irpy", "prev_cent_dist", "spacing_clock", "tally", "level"):
        assert np.array_equal(s2[k], st[k]), k
    assert (s2["step_counter"] == 0).all() and (s2["vel"] == 0).all() and (s2["rpy_rates"] == 0).all()
    env.close()


def test_host_path_equals_device_path():
    E = 128
    a, _ = _make(E)
    b, _ = _make(E)
    a.reset(); b.reset_host()
    rng = np.random.default_rng(0)
    seen = 0
    for t in range(30):
        act = rng.uniform(-1, 1, (E, 3, 4)).astype(np.float32)
        oa, ra, ta, ua = a.step(torch.from_numpy(act).cuda())
        ob, rb, tb, ub, idx, tobs = b.step_host(act)
        assert np.array_equal(oa.cpu().numpy(), ob) and np.array_equal(ra.cpu().numpy(), rb)
        assert np.array_equal(ta.cpu().numpy(), tb) and np.array_equal(ua.cpu().numpy(), ub)
        done = np.nonzero(tb | ub)[0]
        assert sorted(idx.tolist()) == done.tolist()
        for r, e in enumerate(idx):
            assert np.array_equal(tobs[r], a.terminal_obs[e].cpu().numpy())
        seen += len(idx)
    assert seen > 10
    a.close(); b.close()


def test_results_do_not_depend_on_launch_shape_or_sharding():
    """Same env ids => same results for every CTA shape -- including the wave-fitted ones, where a CTA handles a
    non-power-of-two number of envs and leaves the rest of its power-of-two slots empty -- and for 1 shard of 96 envs vs 3
    shards of 32 (the multi-GPU sharding rule, exercised on one GPU)."""
    E = 96
    ref, _ = _make(E, envs_per_block=8, threads_per_block=128)
    variants = [_make(E, envs_per_block=g, threads_per_block=th)[0] for g, th in ((2, 64), (4, 256), (32, 128), (16, 512), (28, 512), (6, 96), (14, 256))]
    # one env per CTA cannot stage the herd with 16-byte async copies and runs the full build of the kernel: another
    # compilation of the same source, equal up to fp32 contraction (like the register variants below)
    other_build = _make(E, envs_per_block=1, threads_per_block=64)[0]
    other_build.reset()
    # the shards get an explicit CTA shape: the default for 32 envs is one env per CTA, i.e. the other build (above)
    shards = [_make(32, env_id_offset=32 * i, envs_per_block=4, threads_per_block=64)[0] for i in range(3)]
    for env in [ref] + variants + shards:
        env.reset()
    gen = torch.Generator("cuda").manual_seed(9)
    for t in range(25):
        act = torch.rand((E, 3, 4), device="cuda", generator=gen) * 2 - 1
        o, r, te, tr = [x.clone() for x in ref.step(act)]
        for v in variants:
            o2, r2, te2, tr2 = v.step(act)
            assert torch.equal(o, o2) and torch.equal(r, r2) and torch.equal(te, te2) and torch.equal(tr, tr2), t
        if t < 3:     # free running: a 1-ulp difference grows about 6x per step under random actions
            o2, r2, te2, tr2 = other_build.step(act)
            assert torch.allclose(o, o2, rtol=1e-4, atol=1e-4) and torch.allclose(r, r2, rtol=1e-4, atol=1e-3)
            assert torch.equal(te, te2) and torch.equal(tr, tr2), t
        for i, sh in enumerate(shards):
            sl = slice(32 * i, 32 * i + 32)
            o2, r2, te2, tr2 = sh.step(act[sl].contiguous())
            assert torch.equal(o[sl], o2) and torch.equal(r[sl], r2) and torch.equal(te[sl], te2) and torch.equal(tr[sl], tr2), (t, i)
    for env in [ref, other_build] + variants + shards:
        env.close()


def test_register_variants_agree():
    """The occupancy build (64 registers) and the latency build (128) are the same arithmetic up to fp32
    contraction differences."""
    E = 64
    a, _ = _make(E, auto_reset=False)
    b, _ = _make(E, auto_reset=False)
    pkg = _pkg()
    if not hasattr(a.lib, "chs_set_register_variant"):
        pytest.skip("library without chs_set_register_variant")
    pkg._lib.check(a.lib.chs_set_register_variant(a._h, 1))
    pkg._lib.check(b.lib.chs_set_register_variant(b._h, 2))
    a.reset(); b.reset()
    act = torch.rand((E, 3, 4), device="cuda", generator=torch.Generator("cuda").manual_seed(1)) * 2 - 1
    oa, ra, _, _ = a.step(act)
    ob, rb, _, _ = b.step(act)
    assert (oa - ob).abs().max().item() <= 2e-6 and (ra - rb).abs().max().item() <= 1e-5
    a.close(); b.close()


@pytest.mark.parametrize("kw", [dict(), dict(aero=7), dict(packing="marl", num_drones=4, fixed_num_drones=4, start_level=2),
                                dict(num_drones=12, fixed_num_drones=0, num_cattle=16, start_level=7)])
def test_split_mode_equals_fused_mode(kw):
    """drone_kernel + step_kernel == the fused kernel (same device functions; fp32 may differ by FMA contraction
    between the two compilations, fp64 and every integer / flag must agree exactly over auto-resets)."""
    for precision in ("fp64", "fp32"):
        E = 200
        a, _ = _make(E, precision=precision, step_mode="fused", **dict(kw))
        b, _ = _make(E, precision=precision, step_mode="split", **dict(kw))
        a.reset(); b.reset()
        D = a.max_drones
        gen = torch.Generator("cuda").manual_seed(5)
        for t in range(30):
            act = torch.rand((E, D, 4), device="cuda", generator=gen) * 2 - 1
            oa, ra, ta, ua = a.step(act)
            ob, rb, tb, ub = b.step(act)
            assert torch.equal(ta, tb) and torch.equal(ua, ub), (precision, t)
            tol = 0.0 if precision == "fp64" else 2e-6
            assert (oa - ob).abs().max().item() <= tol and (ra - rb).abs().max().item() <= tol * 10, (precision, t)
            if precision == "fp32":      # keep the twins on the same trajectory: the roll loop amplifies 1-ulp differences
                b.set_state(a.get_state())
        sa, sb = a.get_state(), b.get_state()
        for k in ("n_drones", "level", "tally", "episode", "step_counter"):
            assert np.array_equal(sa[k], sb[k]), (precision, k)
        a.close(); b.close()


def test_error_paths():
    pkg = _pkg()
    env = pkg.CattleHerdingBatchedEnv(4, num_drones=3, num_cattle=20, device=0)
    with pytest.raises(RuntimeError, match="spawn"):
        env.reset()
    with pytest.raises(RuntimeError, match="cows per simulation"):
        env.set_spawn_table(pkg.spawn.generate(2, 16))       # the reference raises IndexError here
    with pytest.raises(RuntimeError):
        env.set_launch(3, 128)
    env.close()


def test_free_running_fp32_vs_oracle_uniform_actions():
    """fp32 build vs float64 oracle, whole episodes without re-injection under uniform actions: same done flags,
    position divergence bounded (the roll loop amplifies rounding by ~x6 per step, episodes last ~14 steps)."""
    E, M, N = 6, 20, 3
    env, table = _make(E, auto_reset=False, seed=21)
    env.reset()
    oracles = [co.OracleEnv(M, spawn_table=table) for _ in range(E)]

    def sync(idx):
        st = env.get_state()
        for e in idx:
            oracles[e].set_state({k: _pad12(st[k][e]) for k in co.STATE_FIELDS})
    sync(range(E))
    rng = np.random.default_rng(5)
    worst, episodes = 0.0, 0
    for t in range(60):
        a = rng.uniform(-1, 1, (E, N, 4)).astype(np.float32)
        _, _, term, trunc = env.step(torch.from_numpy(a).cuda())
        done = (term | trunc).bool().cpu().numpy()
        st = env.get_state()
        redo = []
        for e in range(E):
            _, _, oterm, otrunc, _ = oracles[e].step(a[e])
            assert bool(done[e]) == bool(oterm or otrunc), (t, e)
            worst = max(worst, float(np.abs(st["pos"][e] - oracles[e].pos[:N]).max()))
            if done[e]:
                redo.append(e)
        if redo:
            m = torch.zeros(E, dtype=torch.uint8); m[redo] = 1
            env.reset(mask=m.cuda()); sync(redo); episodes += len(redo)
    assert episodes >= 10 and worst < 1e-3, (episodes, worst)
    env.close()


def test_autotune_launch_picks_a_shape_and_changes_no_result():
    """`autotune_launch` times a probe env and only changes this env's launch shape: state, counters and statistics are
    untouched, and what it steps afterwards equals an untuned env with the same seeds bit for bit."""
    E = 2048
    tuned, _ = _make(E)
    tuned.reset()
    cands = [(8, 128), (16, 256), (14, 256), (28, 512)]
    best = tuned.autotune_launch(candidates=cands, steps=8)
    assert (best[1], best[2]) in cands and best[0] > 0 and best[3] == "auto"
    assert tuned.dims["envs_per_block"] == best[1] and tuned.dims["threads_per_block"] == best[2]
    assert tuned.stats()[10] == 0
    plain, _ = _make(E, envs_per_block=8, threads_per_block=128)
    plain.reset()
    assert torch.equal(tuned.obs, plain.obs)
    gen = torch.Generator("cuda").manual_seed(3)
    for t in range(6):
        act = torch.rand((E, 3, 4), device="cuda", generator=gen) * 2 - 1
        a, b = tuned.step(act), plain.step(act)
        for x, y in zip(a, b):
            assert torch.equal(x, y), t
    tuned.close(); plain.close()
