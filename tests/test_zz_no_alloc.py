"""GPU: `chs_step_static_scratch` (the allocation-free variant of the split-mode step whose scratch is a static device
array) against the existing entry point `chs_step`, through the C ABI, on two calls in a row.

Collected last on purpose (file name): a failure here hides nothing else.  The test
builds its own envs and does not depend on anything other tests leave behind, so it passes alone in a fresh process and
after the rest of the suite (the static array keeps the previous call's contents; the entry point clears what has to
start at zero).
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

# the fp32 tolerance of tests/test_cuda_parity.py: |err| <= 1e-5 * max(1, |ref|_inf) + 1e-6; flags bit exact
FP32_REL, FP32_ABS = 1e-5, 1e-6


def _make(E, **kw):
    import rl_cattle_herding_b200 as pkg
    M = kw.pop("num_cattle", 20)
    table = pkg.spawn.generate(7, M, seed=0)
    args = dict(num_drones=3, num_cattle=M, fixed_num_drones=3, spawn_table=table, device=0, seed=11, auto_reset=True,
                step_mode="split")
    args.update(kw)
    return pkg.CattleHerdingBatchedEnv(E, **args)


def _close(a, b, what):
    a, b = a.float().cpu().numpy().astype(np.float64), b.float().cpu().numpy().astype(np.float64)
    tol = FP32_REL * max(1.0, float(np.max(np.abs(b)))) + FP32_ABS
    err = float(np.max(np.abs(a - b))) if a.size else 0.0
    assert err <= tol, (what, err, tol)


@pytest.mark.parametrize("kw", [dict(), dict(num_drones=12, fixed_num_drones=0, num_cattle=16), dict(precision="fp64")],
                         ids=["3d20c", "4to12d16c", "fp64"])
def test_static_scratch_step_equals_chs_step_two_calls_in_a_row(kw):
    E = 2048
    ref, new = _make(E, **dict(kw)), _make(E, **dict(kw))
    ref.reset(); new.reset()
    D = ref.max_drones
    gen = torch.Generator("cuda").manual_seed(3)
    done_total = 0
    for call in range(2 + 14):          # the two calls the brief asks for, then on through auto-resets
        act = torch.rand((E, D, 4), device="cuda", generator=gen) * 2 - 1
        o1, r1, t1, u1 = [x.clone() for x in ref.step(act)]                    # existing entry point: chs_step (split mode)
        o2, r2, t2, u2 = [x.clone() for x in new.step_static_scratch(act)]      # new entry point, raw device pointers, same stream
        torch.cuda.synchronize()
        assert torch.equal(t1, t2) and torch.equal(u1, u2), call
        _close(o2, o1, ("obs", call)); _close(r2, r1, ("reward", call))
        done = (t1 | u1).bool()
        if done.any():
            _close(new.terminal_obs[done], ref.terminal_obs[done], ("terminal_obs", call))
        done_total += int(done.sum())
    sa, sb = ref.get_state(), new.get_state()
    for k in ("n_drones", "level", "tally", "episode", "step_counter", "spawn_idx"):
        assert np.array_equal(sa[k], sb[k]), k
    for k in ("pos", "vel", "cattle_vel", "pid_ipos"):
        assert np.allclose(sa[k], sb[k], rtol=FP32_REL, atol=FP32_ABS), k
    assert done_total > 100                      # the finished-env list in the static array was exercised
    assert np.allclose(ref.stats()[:11], new.stats()[:11], rtol=1e-9)     # same episodes ended; float sums in atomic order
    ref.close(); new.close()


def test_static_scratch_is_shared_and_survives_other_handles():
    """Two handles alternate on the one array in stream order (documented as safe), and a plain chs_step in between
    changes nothing."""
    a, b, ref_a, ref_b = _make(512), _make(768, seed=5), _make(512), _make(768, seed=5)
    for e in (a, b, ref_a, ref_b):
        e.reset()
    gen = torch.Generator("cuda").manual_seed(9)
    for call in range(6):
        xa = torch.rand((512, 3, 4), device="cuda", generator=gen) * 2 - 1
        xb = torch.rand((768, 3, 4), device="cuda", generator=gen) * 2 - 1
        oa = [x.clone() for x in (a.step_static_scratch(xa) if call != 3 else a.step(xa))]
        ob = [x.clone() for x in b.step_static_scratch(xb)]
        ra, rb = ref_a.step(xa), ref_b.step(xb)
        for got, want in ((oa, ra), (ob, rb)):
            assert torch.equal(got[2], want[2]) and torch.equal(got[3], want[3]), call
            _close(got[0], want[0], ("obs", call)); _close(got[1], want[1], ("reward", call))
    for e in (a, b, ref_a, ref_b):
        e.close()


def test_sizes_beyond_the_static_array_are_rejected():
    import rl_cattle_herding_b200 as pkg
    E, D = 50000, 12                              # 256 + 4E + 16 D E = 9.8 MB > CHS_STATIC_SCRATCH_BYTES (8 MiB)
    assert 256 + 4 * E + 16 * D * E > pkg._lib.STATIC_SCRATCH_BYTES
    env = _make(E, num_drones=D, fixed_num_drones=4, num_cattle=16)
    env.reset()
    act = torch.zeros((E, D, 4), device="cuda")
    with pytest.raises(RuntimeError, match="CHS_STATIC_SCRATCH_BYTES"):
        env.step_static_scratch(act)
    env.step(act)                                 # the allocating path still serves this batch
    torch.cuda.synchronize()
    env.close()
    small = _make(40000, num_drones=D, fixed_num_drones=4, num_cattle=16)     # 7.84 MB: fits
    small.reset()
    small.step_static_scratch(torch.zeros((40000, D, 4), device="cuda"))
    torch.cuda.synchronize()
    small.close()
    l = pkg._lib.load()
    assert l.chs_step_static_scratch(None, None, None, None, None, None, None, None) == -1
