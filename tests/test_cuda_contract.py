"""GPU: memory-contract checks of the step / reset kernels -- what they may write and that they write it the same way
every time.  compute-sanitizer (memcheck / racecheck) is closed on this GPU pool (profiles/r2g_compute_sanitizer_closed.txt),
so the evidence it would give is collected with checks of our own:

  * guard bands around every caller-owned output tensor stay intact (no out-of-bounds write), for batch sizes that are
    not a multiple of the CTA's env count, fused and split, with auto-reset;
  * the persistent-padding contract: rows >= max_drones and the dead neighbour columns of live rows are written by
    chs_reset only -- a sentinel planted there survives every step (and with the contract off they are rewritten);
  * repeated launches from the same state with the same actions give the same bits (a shared-memory or global-memory race
    between the warps of a CTA -- named barriers, bar.arrive, bulk-async stores, fence.proxy.async in csrc/chs_kernel.cuh --
    would show as run-to-run differences), over every CTA shape the bench uses, fused and split, resets included.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

SENT = 12345.0


def _make(E, **kw):
    import rl_cattle_herding_b200 as pkg
    M = kw.pop("num_cattle", 20)
    args = dict(num_drones=3, num_cattle=M, fixed_num_drones=3, spawn_table=pkg.spawn.generate(9, M, seed=1), device=0, seed=4,
                auto_reset=True)
    args.update(kw)
    return pkg.CattleHerdingBatchedEnv(E, **args)


@pytest.mark.parametrize("mode,E,shape", [("fused", 1003, (8, 128)), ("fused", 517, (16, 256)), ("split", 1003, (32, 256)),
                                          ("fused", 61, (4, 64)), ("split", 2049, (16, 256))])
def test_guard_bands_around_outputs_stay_intact(mode, E, shape):
    env = _make(E, step_mode=mode, envs_per_block=shape[0], threads_per_block=shape[1])
    R, L, gb = env.obs_rows, env.obs_len, 4          # 4 guard envs = 12.7 KB on either side of each image tensor
    dev = env.device
    big_obs = torch.full((E + 2 * gb, R, L), SENT, device=dev)
    big_tobs = torch.full((E + 2 * gb, R, L), SENT, device=dev)
    big_rew = torch.full((E + 32,), SENT, device=dev)
    big_term = torch.full((E + 64,), 77, dtype=torch.uint8, device=dev)
    big_trunc = torch.full((E + 64,), 77, dtype=torch.uint8, device=dev)
    env.obs, env.terminal_obs = big_obs[gb:gb + E], big_tobs[gb:gb + E]
    env.obs.zero_(); env.terminal_obs.zero_()
    env.reward, env.terminated, env.truncated = big_rew[16:16 + E], big_term[32:32 + E], big_trunc[32:32 + E]
    env.reset()
    gen = torch.Generator("cuda").manual_seed(0)
    for _ in range(30):
        env.step(torch.rand((E, 3, 4), device="cuda", generator=gen) * 2 - 1)
    torch.cuda.synchronize()
    assert env.stats()[0] > 0                         # episodes ended: terminal observations and in-kernel resets ran
    for big, lo, hi in ((big_obs, gb, gb + E), (big_tobs, gb, gb + E), (big_rew, 16, 16 + E)):
        assert bool((big[:lo] == SENT).all()) and bool((big[hi:] == SENT).all())
    for big in (big_term, big_trunc):
        assert bool((big[:32] == 77).all()) and bool((big[32 + E:] == 77).all())
        assert set(big[32:32 + E].unique().tolist()) <= {0, 1}
    assert torch.isfinite(env.obs).all() and torch.isfinite(env.reward).all()
    env.close()


@pytest.mark.parametrize("mode", ["fused", "split"])
def test_persistent_padding_is_written_by_reset_only(mode):
    E, D = 512, 3
    env = _make(E, step_mode=mode)
    assert env.persistent_padding
    env.reset()
    # plant a sentinel where the contract says chs_step does not write: the padding rows (the live rows leave in 16-byte
    # units, so the first 12 bytes of row D may be rewritten with zeros: include/chs.h)
    env.obs[:, D:, :] = SENT
    env.obs[:, D, :4] = 0
    env.terminal_obs.zero_()
    gen = torch.Generator("cuda").manual_seed(1)
    resets = torch.zeros(E, dtype=torch.bool, device="cuda")
    for _ in range(25):
        _, _, te, tr = env.step(torch.rand((E, D, 4), device="cuda", generator=gen) * 2 - 1)
        resets |= (te | tr).bool()
    torch.cuda.synchronize()
    pad = env.obs[:, D:, :].clone()
    assert bool((pad[:, 0, :4] == 0).all())
    pad[:, 0, :4] = SENT
    assert bool((pad == SENT).all()) and int(resets.sum()) > 50         # never touched, not even by in-kernel resets
    assert bool((env.obs[:, :D, 14:34] == 0).all())                      # device path: dead columns are (re)written as zeros
    live = torch.cat([env.obs[:, :D, :14], env.obs[:, :D, 34:]], dim=2)
    assert torch.isfinite(live).all() and bool((live[:, :, 0] > 0).all())
    assert bool((env.terminal_obs[:, D:, :] == 0).all())
    env.reset()                                                          # a full reset writes the whole image again
    assert bool((env.obs[:, D:, :] == 0).all())
    env.close()
    # contract off: every step rewrites the full image
    off = _make(E, step_mode=mode, persistent_padding=False)
    off.reset()
    off.obs[:, D:, :] = SENT
    off.step(torch.zeros((E, D, 4), device="cuda"))
    torch.cuda.synchronize()
    assert bool((off.obs[:, D:, :] == 0).all()) and bool((off.obs[:, :D, 14:34] == 0).all())
    off.close()


def test_host_path_skips_padding_and_dead_columns():
    """chs_step_host writes straight into the caller's pinned buffers: padding rows and the 16-byte aligned interior of the
    dead neighbour columns (cols 16..31 lie inside it for every row of a (12, 66) image) are left alone; an in-kernel
    reset rewrites its env's live rows in full; chs_reset_host rewrites everything."""
    E, D = 512, 3
    env = _make(E)
    obs = env.reset_host()
    obs[:, D:, :] = SENT
    obs[:, D, :4] = 0
    obs[:, :D, 16:32] = SENT
    rng = np.random.default_rng(1)
    resets = np.zeros(E, dtype=bool)
    for _ in range(25):
        o, r, te, tr, idx, tobs = env.step_host(rng.uniform(-1, 1, (E, D, 4)).astype(np.float32))
        resets |= (te | tr).astype(bool)
        assert np.all(tobs[:, D:, :] == 0) and np.all(tobs[:, :D, 14:34] == 0)
    assert o is obs or np.shares_memory(o, obs)
    pad = obs[:, D:, :].copy()
    pad[:, 0, :4] = SENT
    assert np.all(pad == SENT)
    dead = obs[:, :D, 16:32]
    assert np.all(dead[~resets] == SENT) and np.all(dead[resets] == 0) and resets.sum() > 50
    obs = env.reset_host()
    assert np.all(obs[:, D:, :] == 0) and np.all(obs[:, :D, 14:34] == 0)
    env.close()


def test_padding_contract_gives_the_same_observations():
    a, b = _make(700, persistent_padding=True), _make(700, persistent_padding=False)
    a.reset(); b.reset()
    gen = torch.Generator("cuda").manual_seed(2)
    for t in range(30):
        act = torch.rand((700, 3, 4), device="cuda", generator=gen) * 2 - 1
        ra, rb = a.step(act), b.step(act)
        assert all(torch.equal(x, y) for x, y in zip(ra, rb)), t
        done = (ra[2] | ra[3]).bool()
        assert torch.equal(a.terminal_obs[done], b.terminal_obs[done])
    a.close(); b.close()


@pytest.mark.parametrize("mode,shape", [("fused", (8, 128)), ("fused", (16, 256)), ("fused", (4, 64)), ("fused", (32, 256)),
                                        ("split", (32, 256)), ("split", (16, 128))])
def test_repeated_launches_are_bit_identical(mode, shape):
    """Race detection by repetition: 12 launches from the same state with the same actions, on a batch in its steady state
    (about 7 % of the envs finish and are rebuilt inside the launch), must agree bit for bit in every output and in the
    complete simulator state."""
    E = 4096
    env = _make(E, step_mode=mode, envs_per_block=shape[0], threads_per_block=shape[1])
    env.reset()
    gen = torch.Generator("cuda").manual_seed(3)
    for _ in range(30):
        env.step(torch.rand((E, 3, 4), device="cuda", generator=gen) * 2 - 1)
    base = env.get_state()
    obs0 = env.obs.clone()
    act = torch.rand((E, 3, 4), device="cuda", generator=gen) * 2 - 1
    ref = None
    for rep in range(12):
        env.set_state(base)
        env.obs.copy_(obs0)
        env.terminal_obs.zero_()
        out = [x.clone() for x in env.step(act)] + [env.terminal_obs.clone()]
        torch.cuda.synchronize()
        st = env.get_state()
        if ref is None:
            ref = (out, st)
            assert int((out[2] | out[3]).sum()) > 100
            continue
        for x, y in zip(out, ref[0]):
            assert torch.equal(x, y), (mode, shape, rep)
        for k in st:
            assert np.array_equal(st[k], ref[1][k], equal_nan=True), (mode, shape, rep, k)
    env.close()


def test_host_path_counts_and_terminal_rows_over_many_calls():
    """The host path's finished-env counter runs in ping-pong between two device words; mixing calls with and without the
    terminal-observation outputs must not leave a stale count behind."""
    import ctypes as C
    import rl_cattle_herding_b200 as pkg
    E = 640
    env, twin = _make(E), _make(E)
    env.reset_host(); twin.reset()
    rng = np.random.default_rng(0)
    seen = 0
    for t in range(40):
        act = rng.uniform(-1, 1, (E, 3, 4)).astype(np.float32)
        ot, rt, tt, ut = twin.step(torch.from_numpy(act).cuda())
        if t % 5 == 3:
            # a call without the terminal-observation outputs (NULL pointers) through the raw ABI
            hb = env._host_buffers()
            np.copyto(hb["actions"], act)
            h = env._host
            pkg._lib.check(env.lib.chs_step_host(env._h, C.c_void_p(h["actions"].data_ptr()), C.c_void_p(h["obs"].data_ptr()),
                                                 C.c_void_p(h["reward"].data_ptr()), C.c_void_p(h["terminated"].data_ptr()),
                                                 C.c_void_p(h["truncated"].data_ptr()), None, None, None, env._stream()))
            obs, term, trunc = hb["obs"], hb["terminated"], hb["truncated"]
        else:
            obs, rew, term, trunc, idx, tobs = env.step_host(act)
            done = np.nonzero(term | trunc)[0]
            assert sorted(idx.tolist()) == done.tolist(), t
            for r, e in enumerate(idx):
                assert np.array_equal(tobs[r], twin.terminal_obs[e].cpu().numpy())
            seen += len(idx)
        assert np.array_equal(obs, ot.cpu().numpy()) and np.array_equal(term, tt.cpu().numpy()) and np.array_equal(trunc, ut.cpu().numpy())
    assert seen > 100
    env.close(); twin.close()


@pytest.mark.parametrize("knob", ["CHS_NO_PDL", "CHS_HOST_COPY", "CHS_NO_DIRECT", "CHS_EARLY_ROWS", "CHS_DEFERRED_RESET", "CHS_NO_TMA"])
def test_launch_and_host_path_knobs_do_not_change_results(knob, monkeypatch):
    """The step kernels are launched with programmatic dependent launch and the pinned host path is copy free; the knobs
    that turn either off (plain launches / results through the staging copy / everything staged in HBM), and the one that
    moves split mode's rebuild of finished envs into a third kernel and the one that stages the herd with cp.async chunks
    instead of TMA tile loads, are read at chs_create and must give the same bits: device path against device path, host path against host path, 30 steps with
    resets, fused and split."""
    E = 700
    rng = np.random.default_rng(5)
    acts = [rng.uniform(-1, 1, (E, 3, 4)).astype(np.float32) for _ in range(30)]

    def run(mode):
        env = _make(E, step_mode=mode)
        env.reset_host()
        host, dev = [], []
        for a in acts:
            obs, rew, term, trunc, idx, tobs = env.step_host(a)
            host.append((obs.copy(), rew.copy(), term.copy(), trunc.copy(), np.sort(idx.copy()),
                         tobs[np.argsort(idx)].copy()))
        twin = _make(E, step_mode=mode)
        twin.reset()
        for a in acts:
            out = twin.step(torch.from_numpy(a).cuda())
            dev.append(tuple(x.clone() for x in out) + (twin.terminal_obs.clone(),))
        st = twin.get_state()
        env.close(); twin.close()
        return host, dev, st

    for mode in ("fused", "split"):
        monkeypatch.delenv(knob, raising=False)
        ref = run(mode)
        monkeypatch.setenv(knob, "1")
        got = run(mode)
        for t, (a, b) in enumerate(zip(ref[0], got[0])):
            for x, y in zip(a, b):
                assert np.array_equal(x, y), (knob, mode, "host", t)
        for t, (a, b) in enumerate(zip(ref[1], got[1])):
            for x, y in zip(a, b):
                assert torch.equal(x, y), (knob, mode, "device", t)
        for k in ref[2]:
            assert np.array_equal(ref[2][k], got[2][k], equal_nan=True), (knob, mode, k)
