"""Learner-side checks: the PPO restatement against a plain-numpy reading of SB3's formulas (CPU), and short training
runs on the GPU simulator for both packings."""
import math
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import rl_cattle_herding_b200 as pkg  # noqa: E402
from rl_cattle_herding_b200 import ppo  # noqa: E402


def test_gae_matches_sb3_loop():
    """RolloutBuffer.compute_returns_and_advantage written out in numpy (stable_baselines3/common/buffers.py)."""
    rng = np.random.default_rng(0)
    T, B, gamma, lam = 7, 5, 0.99, 0.95
    rew, val = rng.normal(size=(T, B)), rng.normal(size=(T, B))
    done = (rng.random((T, B)) < 0.3).astype(np.float64)
    last = rng.normal(size=(B,))
    adv = np.zeros((T, B))
    gae_ = np.zeros(B)
    for t in reversed(range(T)):
        if t == T - 1:
            nonterm, nv = 1.0 - done[t], last       # SB3 shifts: episode_starts[t+1] == done[t]
        else:
            nonterm, nv = 1.0 - done[t], val[t + 1]
        delta = rew[t] + gamma * nv * nonterm - val[t]
        gae_ = delta + gamma * lam * nonterm * gae_
        adv[t] = gae_
    a, r = ppo.gae(torch.tensor(rew), torch.tensor(val), torch.tensor(done), torch.tensor(last), gamma, lam)
    assert np.allclose(a.numpy(), adv, atol=1e-12) and np.allclose(r.numpy(), adv + val, atol=1e-12)


def test_policy_shapes_and_distribution():
    torch.manual_seed(0)
    pol = ppo.MlpActorCritic(792, 12, (128, 128), -1.0)
    n_par = sum(p.numel() for p in pol.parameters())
    # two towers 792-128-128, heads 128->12 and 128->1, log_std 12  (SB3 MlpPolicy with this net_arch)
    assert n_par == 2 * (792 * 128 + 128 + 128 * 128 + 128) + (128 * 12 + 12) + (128 + 1) + 12
    obs = torch.randn(9, 792)
    a, logp, v = pol.act(obs)
    assert a.shape == (9, 12) and logp.shape == (9,) and v.shape == (9,)
    mean, log_std = pol.dist(obs)
    ref = torch.distributions.Normal(mean, log_std.exp().expand_as(mean))
    assert torch.allclose(pol.log_prob(mean, log_std, a), ref.log_prob(a).sum(-1), atol=1e-5)
    assert math.isclose(float(pol.entropy(log_std)), float(ref.entropy().sum(-1)[0]), rel_tol=1e-6)
    assert torch.allclose(log_std, torch.full((12,), -1.0))


def test_loss_equals_sb3_formula_and_ignores_padding_samples():
    """All weights 1: SB3's `PPO.train` loss written out.  Weight 0 (MARL slots >= NUM_DRONES): the sample's contents,
    NaN included, change neither the loss nor any gradient (phantom agents entered the loss)."""
    torch.manual_seed(1)
    pol = ppo.MlpActorCritic(66, 4, (32, 32), 0.0).double()
    hp = dict(ppo.RLLIB_DTDE)
    n = 40
    obs, act = torch.randn(n, 66, dtype=torch.float64), torch.randn(n, 4, dtype=torch.float64)
    old, adv, ret = torch.randn(n, dtype=torch.float64) - 4, torch.randn(n, dtype=torch.float64), torch.randn(n, dtype=torch.float64)
    loss, pg, vf, ent, cf = ppo.ppo_loss(pol, obs, act, old, adv, ret, torch.ones(n, dtype=torch.float64), hp)
    a = (adv - adv.mean()) / (adv.std() + 1e-8)
    mu, ls = pol.dist(obs)
    ratio = torch.exp(pol.log_prob(mu, ls, act) - old)
    pg_ref = -torch.min(a * ratio, a * ratio.clamp(1 - hp["clip_range"], 1 + hp["clip_range"])).mean()
    vf_ref = torch.nn.functional.mse_loss(ret, pol.value(obs))
    assert torch.allclose(pg, pg_ref, atol=1e-12) and torch.allclose(vf, vf_ref, atol=1e-12)
    assert torch.allclose(loss, pg_ref + hp["ent_coef"] * (-pol.entropy(ls).mean()) + hp["vf_coef"] * vf_ref, atol=1e-12)
    # 15 padding samples with garbage spliced in between the real ones
    m = n + 15
    keep = torch.randperm(m)[:n].sort().values
    w = torch.zeros(m, dtype=torch.float64); w[keep] = 1

    def pad(x, fill):
        out = torch.full((m,) + x.shape[1:], fill, dtype=torch.float64)
        out[keep] = x
        return out
    grads = []
    for fill in (0.0, 1e6, float("nan")):
        pol.zero_grad()
        l2, pg2, vf2, _, cf2 = ppo.ppo_loss(pol, pad(obs, fill), pad(act, fill), pad(old, fill), pad(adv, fill), pad(ret, fill), w, hp)
        l2.backward()
        assert torch.allclose(l2, loss, atol=1e-12) and torch.allclose(pg2, pg, atol=1e-12) and torch.allclose(cf2, cf, atol=1e-12), fill
        grads.append([p.grad.clone() for p in pol.parameters()])
    pol.zero_grad()
    loss.backward()
    for g in grads:
        for a_, b_ in zip(g, pol.parameters()):
            assert torch.isfinite(a_).all() and torch.allclose(a_, b_.grad, atol=1e-10)


def test_reference_hyperparameters():
    """simulator/CTDECattleHerder.py:107-127 and simulator/DTDECattleHerder.py:84-93."""
    h = ppo.SB3_CTDE
    assert (h["learning_rate"], h["n_epochs"], h["gamma"], h["gae_lambda"], h["clip_range"], h["ent_coef"], h["vf_coef"],
            h["max_grad_norm"], h["hidden"], h["log_std_init"]) == (3e-4, 10, 0.99, 0.95, 0.1, 0.1, 0.7, 0.5, (128, 128), -1.0)
    r = ppo.RLLIB_DTDE
    assert (r["learning_rate"], r["n_epochs"], r["gamma"], r["clip_range"], r["ent_coef"], r["vf_coef"]) == (3e-4, 10, 0.99, 0.1, 0.01, 0.7)


@pytest.mark.gpu
@pytest.mark.parametrize("packing", ["ctde", "marl"])
def test_short_training_run(packing):
    E, N, M = 256, 3, 20
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, ctrl_freq=30, fixed_num_drones=N,
                                      start_level=7 if packing == "ctde" else 0, spawn_table=pkg.spawn.generate(16, M, seed=0),
                                      device=0, seed=0, packing=packing)
    algo = ppo.DevicePPO(env, n_steps=8, seed=0, n_epochs=2, cuda_graphs=(packing == "ctde"))   # one graph run, one eager
    assert algo.B == (E if packing == "ctde" else E * N)
    assert algo.obs_dim == (12 * 66 if packing == "ctde" else 66) and algo.act_dim == (N * 4 if packing == "ctde" else 4)
    before = [p.detach().clone() for p in algo.policy.parameters()]
    hist = algo.learn(2)
    assert len(hist) == 2 and algo.num_timesteps == 2 * 8 * E
    for h in hist:
        assert all(math.isfinite(h[k]) for k in ("pg", "vf", "ent", "clipfrac", "explained_variance", "env_steps_per_s"))
    assert any(not torch.equal(a, b.detach()) for a, b in zip(before, algo.policy.parameters()))
    # the rollout buffer holds what the simulator produced: rewards finite, done flags 0/1
    assert torch.isfinite(algo.buf["rew"]).all() and set(algo.buf["done"].unique().tolist()) <= {0.0, 1.0}
    # done flags in the last rollout = episodes the simulator's own accumulator counted during it
    per_env = algo.buf["done"].sum() / (N if packing == "marl" else 1)
    assert float(per_env) == hist[-1]["episodes"]
    env.close()


@pytest.mark.gpu
def test_marl_padding_slots_are_not_samples():
    """4 drone slots, level 0 => 3 drones per episode: slot 3 of every env is padding.  Its weight is 0 in the rollout
    buffer and poisoning its stored observation / advantage changes nothing in the update."""
    E, D, M = 128, 4, 20
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=D, num_cattle=M, ctrl_freq=30, start_level=0, packing="marl",
                                      spawn_table=pkg.spawn.generate(16, M, seed=0), device=0, seed=0)
    algo = ppo.DevicePPO(env, n_steps=6, seed=0, n_epochs=1, cuda_graphs=False)
    algo.collect()
    w = algo.buf["w"].reshape(6, E, D)
    assert torch.equal(w[:, :, :3], torch.ones_like(w[:, :, :3])) and torch.equal(w[:, :, 3], torch.zeros_like(w[:, :, 3]))
    assert torch.isfinite(algo.buf["rew"]).all()
    state = [p.detach().clone() for p in algo.policy.parameters()]
    torch.manual_seed(5)
    algo.update()
    after_clean = [p.detach().clone() for p in algo.policy.parameters()]
    with torch.no_grad():
        for p, q in zip(algo.policy.parameters(), state):
            p.copy_(q)
        for st in algo.opt.state.values():
            for v in st.values():
                if torch.is_tensor(v):
                    v.zero_()
    algo.buf["obs"].reshape(6, E, D, -1)[:, :, 3] = float("nan")
    algo.buf["adv"].reshape(6, E, D)[:, :, 3] = 1e9
    algo.buf["ret"].reshape(6, E, D)[:, :, 3] = -1e9
    torch.manual_seed(5)
    algo.update()
    for a, b in zip(after_clean, algo.policy.parameters()):
        assert torch.isfinite(b).all() and torch.allclose(a, b.detach(), atol=1e-6)
    env.close()
