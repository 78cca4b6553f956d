"""GPU: the oracle on the shapes the performance numbers are quoted on ("What's weak" #1).

`tests/test_cuda_parity.py` lands every recorded step of the reference on the device in small batches with the
default launch.  Here the same checker runs INSIDE the batches bench.py measures -- 4096 envs with every CTA shape the
bench's autotuner may pick, fused and split; 65 536 and 262 144 envs in `step_mode="auto"` (split: drone_kernel +
step_kernel<SPLIT> + reset_list_kernel, 64-register build); BASELINE config 3 (8 drones / 100 cattle x 1024 envs) --
with auto-reset on and the alpha-mask table active: a strided subset of envs is advanced by `OracleEnv` from
`chs_get_state()` on every step (oracle/subset_check.py: flags exact, obs / reward at the fp32 tolerance, first
observation of auto-reset envs against the oracle's reset fed with the device's Philox draws), and the first 64 envs of
a 4096-env batch replay states recorded from the reference itself.
"""
import numpy as np
import pytest

from _golden import FLOAT_STATE, INT_STATE, Golden, max_err

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

import subset_check as sc  # noqa: E402  (tests/_golden.py put oracle/ on sys.path)

# bench.py's candidate list for the 4096-env headline (run_cuda_arm: `cands`)
# (28, 512), (28, 256): wave-fitted shapes -- a CTA handles 28 of its 32 env slots, 147 CTAs on the 148 SMs at 4096 envs
BENCH_SHAPES = [(4, 64), (8, 96), (32, 512), (28, 512), (28, 256)] + [(g, th) for g in (4, 8, 16, 32) for th in (128, 256)]


def _make(E, N=3, M=20, rows=100, **kw):
    import rl_cattle_herding_b200 as pkg
    table = pkg.spawn.generate(rows, M, seed=0)
    args = dict(num_drones=N, num_cattle=M, ctrl_freq=30, fixed_num_drones=N, start_level=7, auto_reset=True,
                spawn_table=table, device=0, seed=0)
    args.update(kw)
    return pkg.CattleHerdingBatchedEnv(E, **args), table


def _run(env, table, ids, steps, N, seed=0, gen_seed=1, fixed=True, warm=0):
    chk = sc.SubsetChecker(env, table, ids, seed=seed, fixed_num_drones=N if fixed else 0)
    gen = torch.Generator("cuda").manual_seed(gen_seed)
    for _ in range(warm):      # into the steady state: episodes of every age, ~7 % of the envs finishing per step
        env.step(torch.rand((env.num_envs, env.max_drones, 4), device="cuda", generator=gen) * 2 - 1)
    for _ in range(steps):
        act = torch.rand((env.num_envs, env.max_drones, 4), device="cuda", generator=gen) * 2 - 1
        chk.step(act)
    return chk


@pytest.mark.parametrize("mode", ["fused", "split"])
@pytest.mark.parametrize("shape", BENCH_SHAPES, ids=lambda s: "G%d_T%d" % s)
def test_4096_envs_every_bench_launch_shape(shape, mode):
    """Headline config, each (envs per CTA, threads per CTA) x step mode the bench may time: 6 steps x 64 envs."""
    env, table = _make(4096, envs_per_block=shape[0], threads_per_block=shape[1], step_mode=mode)
    env.reset()
    chk = _run(env, table, sc.subset_ids(4096, 16, 4), 6, 3, warm=10)
    s = chk.summary()
    assert s["ok"] and s["flags_equal"], (shape, mode, chk.failures[:5], s)
    assert s["env_steps"] == 6 * 64
    env.close()


@pytest.mark.parametrize("E", [4096, 24576, 65536, 262144])
def test_sweep_points_auto_mode_subset_vs_oracle(E):
    """The sweep points of bench.py in step_mode="auto" (fused up to 22 100 envs at 3 drones, split beyond): 16 steps,
    256 envs on every step, resets included."""
    env, table = _make(E)
    env.reset()
    chk = _run(env, table, sc.subset_ids(E, 32, 8), 16, 3)
    s = chk.summary()
    assert s["ok"] and s["flags_equal"], (E, chk.failures[:5], s)
    assert s["env_steps"] == 16 * 256 and s["resets_checked"] >= 50, s
    st = env.stats()
    assert st[10] == 16 * E and st[9] == 0
    env.close()


@pytest.mark.parametrize("variant", [1, 2])
@pytest.mark.parametrize("mode", ["fused", "split"])
@pytest.mark.parametrize("G", [16, 28])
def test_register_variants_and_modes_vs_oracle(variant, mode, G):
    """Both register budgets of the fp32 kernel in both step modes, 8192 envs, against the oracle (not a sibling); G = 28
    is the wave-fitted shape of the large batches (a ragged last CTA: 8192 = 292 x 28 + 16)."""
    import rl_cattle_herding_b200 as pkg
    env, table = _make(8192, step_mode=mode, envs_per_block=G, threads_per_block=256)
    pkg._lib.check(env.lib.chs_set_register_variant(env._h, variant))
    env.reset()
    chk = _run(env, table, sc.subset_ids(8192, 16, 8), 12, 3, warm=10)
    s = chk.summary()
    assert s["ok"] and s["flags_equal"] and s["resets_checked"] > 10, (variant, mode, chk.failures[:5], s)
    env.close()


def test_variable_drone_count_large_batch():
    """Level 7 draws 4..12 drones per episode (BaseAviary.py:307): 16 384 envs x 12 slots, split mode by size."""
    env, table = _make(16384, N=12, M=16, fixed_num_drones=0)
    env.reset()
    chk = _run(env, table, sc.subset_ids(16384, 16, 4), 10, 12, fixed=False, warm=10)
    s = chk.summary()
    assert s["ok"] and s["flags_equal"] and s["resets_checked"] > 5, (chk.failures[:5], s)
    env.close()


@pytest.mark.parametrize("mode", ["fused", "split"])
def test_config3_8d100c_1024_envs(mode):
    """BASELINE config 3 (8 drones / 100 cattle x 1024 envs) against the oracle: 6 steps x 48 envs."""
    env, table = _make(1024, N=8, M=100, rows=8, step_mode=mode)
    env.reset()
    chk = _run(env, table, sc.subset_ids(1024, 12, 4), 6, 8, warm=8)
    s = chk.summary()
    assert s["ok"] and s["flags_equal"], (mode, chk.failures[:5], s)
    env.close()


@pytest.mark.parametrize("name,E,mode,variant", [("a_3d20c_uniform", 4096, "fused", 1), ("a_3d20c_uniform", 4096, "split", 2),
                                                   ("a_3d20c_uniform", 65536, "auto", 0), ("c_8d100c_uniform", 1024, "auto", 0)])
def test_reference_states_replayed_inside_the_batch(name, E, mode, variant):
    """SURVEY config 2: the first 64 envs of the batch replay states RECORDED FROM THE REFERENCE (tests/golden), the
    other envs run their own episodes; one launch steps them all and the 64 must land on the reference's own results."""
    import rl_cattle_herding_b200 as pkg
    g = Golden(name)
    N, M = g.meta["n_drones"], g.meta["n_cattle"]
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, ctrl_freq=g.meta["ctrl_freq"], fixed_num_drones=N,
                                      start_level=g.meta["start_level"], auto_reset=False, spawn_table=g.table, device=0,
                                      seed=0, step_mode=mode)
    pkg._lib.check(env.lib.chs_set_register_variant(env._h, variant))
    env.reset()
    for _ in range(5):
        env.step(torch.rand((E, N, 4), device="cuda") * 2 - 1)
    n = min(64, g.steps)
    for t0 in (0, g.steps - n):
        sl = slice(t0, t0 + n)
        st = {}
        for k in FLOAT_STATE + INT_STATE:
            a = g.z[f"before_{k}"][sl]
            st[k] = a[:, :N] if (a.ndim == 3 and a.shape[1] == 12) else a
        env.set_state(st)
        act = torch.rand((E, N, 4), device="cuda") * 2 - 1
        act[:n] = torch.from_numpy(np.ascontiguousarray(g.z["action"][sl][:, :N])).cuda()
        obs, rew, term, trunc = env.step(act)
        torch.cuda.synchronize()
        got = env.get_state(0, n)
        assert np.array_equal(term[:n].cpu().numpy(), g.z["terminated"][sl]) and np.array_equal(trunc[:n].cpu().numpy(), g.z["truncated"][sl])
        for k in INT_STATE:
            assert np.array_equal(got[k].astype(np.int64), g.z[f"after_{k}"][sl].astype(np.int64)), (name, k)
        for k in FLOAT_STATE:
            ref = g.z[f"after_{k}"][sl]
            ref = ref[:, :N] if (ref.ndim == 3 and ref.shape[1] == 12) else ref
            scale = max(1.0, float(np.max(np.abs(np.where(np.isfinite(ref), ref, 0.0)))))
            tol = 1e-5 * scale + (4e-6 if k == "cattle_vel" else 1e-6)
            assert max_err(got[k], ref) <= tol, (name, E, mode, k, max_err(got[k], ref), tol)
        o = obs[:n].cpu().numpy()
        ref_obs = g.z["obs"][sl]
        for i in range(n):      # neighbour columns of near-tie rows: any equidistant choice (tests/_golden.py)
            o[i] = sc._mask_ties(o[i], ref_obs[i], g.z["after_pos"][t0 + i], N)
        scale = max(1.0, float(np.max(np.abs(ref_obs))))
        assert max_err(o, ref_obs) <= 1e-5 * scale + 1e-6, (name, E, mode, max_err(o, ref_obs))
        ref_rew = g.z["reward"][sl]
        assert max_err(rew[:n].cpu().numpy(), ref_rew) <= 5e-5 * max(1.0, float(np.max(np.abs(ref_rew))))
    env.close()
