"""Device-resident PPO over the batched simulator: the training loops of the reference with the rollout on the GPU.

What it replaces (BASELINE config 5):

* CTDE -- `simulator/CTDECattleHerder.py:91-127`: SB3 `PPO("MlpPolicy", make_vec_env(CattleAviary, n_envs=24,
  vec_env_cls=SubprocVecEnv), learning_rate=3e-4, n_steps=2048, batch_size=64, n_epochs=10, gamma=0.99,
  gae_lambda=0.95, clip_range=0.1, ent_coef=0.1, vf_coef=0.7, max_grad_norm=0.5, policy_kwargs=dict(log_std_init=-1,
  ortho_init=False, net_arch=[dict(pi=[128, 128], vf=[128, 128])]))`.  One policy sees the whole `(12, L)`
  observation and emits all `(num_drones, 4)` actions.
* DTDE -- `simulator/DTDECattleHerder.py:62-100`: RLlib PPO with one shared policy over the `agent_<i>` streams of
  `RLlibMultiAgentWrapper` (gamma 0.99, lr 3e-4, num_sgd_iter 10, vf_loss_coeff 0.7, entropy_coeff 0.01,
  clip_param 0.1).  Every drone is a sample: observation `(L,)`, action `(4,)`, its own reward.

Neither stable_baselines3 nor ray is installed on the build or GPU image, so the learner here is a plain PyTorch
restatement of SB3's on-policy loop (`OnPolicyAlgorithm.collect_rollouts`, `PPO.train`,
`RolloutBuffer.compute_returns_and_advantage`): same loss, same advantage normalisation, same bootstrap of
time-limit truncations with the value of the terminal observation.  What differs is where the data lives: the
simulator writes observations, rewards and flags straight into device tensors, the policy reads them there, the
rollout buffer is a set of `(T, B, ...)` CUDA tensors, and nothing crosses PCIe during an iteration.  The learner's
matrix products go through cuBLAS (library code, not the hot path of this repository -- that is `chs_step`).

With `torch.distributed` initialised (one process per GPU, envs sharded by `dist.shard_range`) gradients are averaged
with one all-reduce per minibatch: per-GPU learners, NCCL over NVLink.
"""
import math
import time

import torch
from torch import nn

SB3_CTDE = dict(learning_rate=3e-4, n_epochs=10, gamma=0.99, gae_lambda=0.95, clip_range=0.1, ent_coef=0.1, vf_coef=0.7,
                max_grad_norm=0.5, hidden=(128, 128), log_std_init=-1.0)
# RLlib PPOConfig defaults where DTDECattleHerder.py does not override them: lambda_ 1.0, fcnet_hiddens [256, 256] tanh,
# grad_clip None.  The adaptive KL penalty of RLlib's PPO (kl_coeff 0.2) is not reproduced: the clipped surrogate alone.
RLLIB_DTDE = dict(learning_rate=3e-4, n_epochs=10, gamma=0.99, gae_lambda=1.0, clip_range=0.1, ent_coef=0.01, vf_coef=0.7,
                  max_grad_norm=None, hidden=(256, 256), log_std_init=0.0)


class MlpActorCritic(nn.Module):
    """SB3 `ActorCriticPolicy` with `net_arch=dict(pi=[...], vf=[...])`, tanh, `ortho_init=False`: two separate towers,
    a linear action head, a linear value head and a state-independent log-std (DiagGaussianDistribution)."""

    def __init__(self, obs_dim, act_dim, hidden=(128, 128), log_std_init=-1.0):
        super().__init__()

        def tower():
            layers, d = [], obs_dim
            for h in hidden:
                layers += [nn.Linear(d, h), nn.Tanh()]
                d = h
            return nn.Sequential(*layers), d

        self.pi, dp = tower()
        self.vf, dv = tower()
        self.action_net = nn.Linear(dp, act_dim)
        self.value_net = nn.Linear(dv, 1)
        self.log_std = nn.Parameter(torch.full((act_dim,), float(log_std_init)))

    def value(self, obs):
        return self.value_net(self.vf(obs)).squeeze(-1)

    def dist(self, obs):
        return self.action_net(self.pi(obs)), self.log_std

    @staticmethod
    def log_prob(mean, log_std, act):
        var = torch.exp(2.0 * log_std)
        return (-0.5 * (act - mean) ** 2 / var - log_std - 0.5 * math.log(2.0 * math.pi)).sum(-1)

    @staticmethod
    def entropy(log_std):
        return (0.5 + 0.5 * math.log(2.0 * math.pi) + log_std).sum(-1)

    @torch.no_grad()
    def act(self, obs):
        mean, log_std = self.dist(obs)
        noise = torch.randn(mean.shape, device=mean.device, dtype=mean.dtype)      # default generator: graph-capture safe
        a = mean + torch.exp(log_std) * noise
        return a, self.log_prob(mean, log_std, a), self.value(obs)


def gae(rewards, values, dones, last_values, gamma, lam):
    """`RolloutBuffer.compute_returns_and_advantage` on `(T, B)` tensors.  `dones[t]` = the episode of sample t ended
    with that step (so the value after it does not belong to it).  Returns (advantages, returns)."""
    T = rewards.shape[0]
    adv = torch.zeros_like(rewards)
    last = torch.zeros_like(last_values)
    nxt = last_values
    for t in range(T - 1, -1, -1):
        nonterminal = 1.0 - dones[t]
        delta = rewards[t] + gamma * nxt * nonterminal - values[t]
        last = delta + gamma * lam * nonterminal * last
        adv[t] = last
        nxt = values[t]
    return adv, adv + values


def ppo_loss(policy, obs, act, old_logp, adv, ret, w, hp):
    """SB3 `PPO.train` loss of one minibatch with per-sample weights `w` (1 = a sample that exists, 0 = padding).
    With all weights 1 this is SB3's loss exactly (per-minibatch advantage normalisation with the unbiased std, clipped
    surrogate, MSE value loss, entropy bonus); a sample with weight 0 contributes to no mean, to no normalisation
    statistic and to no gradient, whatever its contents (they are replaced before any arithmetic, so NaN padding is
    harmless).  Returns (loss, pg, vf, ent, clipfrac)."""
    live = w > 0
    n = w.sum().clamp_min(1.0)
    adv = torch.where(live, adv, torch.zeros_like(adv))
    ret = torch.where(live, ret, torch.zeros_like(ret))
    old_logp = torch.where(live, old_logp, torch.zeros_like(old_logp))
    obs = torch.where(live[:, None], obs, torch.zeros_like(obs))
    act = torch.where(live[:, None], act, torch.zeros_like(act))
    if adv.numel() > 1:
        mean = (adv * w).sum() / n
        var = (((adv - mean) ** 2) * w).sum() / (n - 1.0).clamp_min(1.0)
        adv = (adv - mean) / (var.sqrt() + 1e-8)
    mu, log_std = policy.dist(obs)
    logp = policy.log_prob(mu, log_std, act)
    ratio = torch.exp(logp - old_logp)
    surr = torch.min(adv * ratio, adv * ratio.clamp(1 - hp["clip_range"], 1 + hp["clip_range"]))
    pg = -(surr * w).sum() / n
    vf = (((ret - policy.value(obs)) ** 2) * w).sum() / n
    ent = -policy.entropy(log_std).mean()              # state-independent log-std: the same for every sample
    loss = pg + hp["ent_coef"] * ent + hp["vf_coef"] * vf
    clipfrac = ((((ratio.detach() - 1).abs() > hp["clip_range"]).float()) * w).sum() / n
    return loss, pg, vf, ent, clipfrac


def average_gradients(params):
    """One all-reduce of the flattened gradients across the process group (no-op without torch.distributed)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return
    grads = [p.grad for p in params if p.grad is not None]
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat)
    flat /= dist.get_world_size()
    o = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[o:o + n].view_as(g))
        o += n


class DevicePPO:
    """On-policy PPO whose rollout never leaves the GPU.

    `env` is a `CattleHerdingBatchedEnv` (`packing="ctde"`: one sample per env) or a `GpuMARLBatchedEnv`
    (`packing="marl"`: one sample per drone slot, shared policy).  `n_steps` control steps of all envs make one
    rollout of `n_steps * B` samples; `batch_size` is the minibatch (SB3's 64 is a CPU-sized choice; the default here
    keeps `n_steps * B / batch_size` = 16 minibatches per epoch).

    `cuda_graphs=True` (default on a single GPU) captures the whole rollout -- `n_steps` x (policy, `chs_step`,
    bootstrap value, buffer writes) plus the advantage recursion -- as ONE CUDA graph and one minibatch update
    (forward, backward, gradient clipping, Adam) as another: an iteration is then 1 + n_epochs * minibatches graph
    launches instead of thousands of kernel launches from Python.  `chs_step` is capturable by contract: no
    allocation and no host synchronisation inside, work goes to the stream it is handed."""

    def __init__(self, env, n_steps=16, batch_size=None, seed=0, cuda_graphs=None, **hp):
        self.env = env
        self.marl = env.packing == "marl"
        base = dict(RLLIB_DTDE if self.marl else SB3_CTDE)
        base.update(hp)
        self.hp = base
        E, D = env.num_envs, env.max_drones
        self.B = E * D if self.marl else E
        self.obs_dim = env.obs_len if self.marl else env.obs_rows * env.obs_len
        self.act_dim = env.action_width if self.marl else D * env.action_width
        self.n_steps = int(n_steps)
        self.batch_size = int(batch_size) if batch_size else max(64, self.n_steps * self.B // 16)
        dev = env.device
        torch.manual_seed(seed)
        torch.cuda.manual_seed(seed)
        self.policy = MlpActorCritic(self.obs_dim, self.act_dim, base["hidden"], base["log_std_init"]).to(dev)
        import torch.distributed as dist
        multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        # minibatch update as CUDA graphs.  One GPU: one graph per minibatch step.  Per-GPU learners: TWO graphs with the
        # NCCL gradient all-reduce launched eagerly in between (capturing the collective inside the graph hung on 2 x B200
        # with this torch / NCCL pair): forward + backward into one flat gradient buffer | all-reduce | clip + Adam
        self.multi = multi
        self.cuda_graphs = True if cuda_graphs is None else bool(cuda_graphs)
        # the rollout has no collective in it (envs are sharded, nothing is exchanged): its graph is safe on every rank
        self.graph_rollout = True if cuda_graphs is None else bool(cuda_graphs)
        self.opt = torch.optim.Adam(self.policy.parameters(), lr=base["learning_rate"], eps=1e-5, capturable=self.cuda_graphs)
        T, B = self.n_steps, self.B
        f = dict(dtype=torch.float32, device=dev)
        self.buf = dict(obs=torch.zeros((T, B, self.obs_dim), **f), act=torch.zeros((T, B, self.act_dim), **f),
                        logp=torch.zeros((T, B), **f), val=torch.zeros((T, B), **f), rew=torch.zeros((T, B), **f),
                        done=torch.zeros((T, B), **f), adv=torch.zeros((T, B), **f), ret=torch.zeros((T, B), **f),
                        w=torch.ones((T, B), **f))      # 1 = the sample's drone slot exists (MARL: slot < NUM_DRONES of its env)
        self._slot = torch.arange(D, dtype=torch.int32, device=dev)
        self._idx = torch.zeros((self.batch_size,), dtype=torch.long, device=dev)
        self._acc = torch.zeros((4,), **f)            # sums of pg, vf, ent, clipfrac over the minibatches of an update
        self.num_timesteps = 0
        self._obs = None
        self._g_roll = self._g_upd = self._g_upd_b = None
        self._flat = None
        self.last = {}

    # ---- views of the simulator's outputs as (B, ...) samples --------------------------------------------------
    def _flat_obs(self, obs):
        return obs.reshape(self.B, self.obs_dim)

    def _sample_flags(self):
        """(done, bootstrap) per sample, float: the sample's episode ended / ended by truncation only."""
        env = self.env
        if self.marl:
            D = env.max_drones
            term_all, trunc_all = env.terminated[:, D].bool(), env.truncated[:, D].bool()
            done = (term_all | trunc_all)
            boot = done & ~term_all
            return (done[:, None].expand(-1, D).reshape(-1).float(), boot[:, None].expand(-1, D).reshape(-1).float())
        term, trunc = env.terminated.bool(), env.truncated.bool()
        return (term | trunc).float(), (trunc & ~term).float()

    # ---- rollout -----------------------------------------------------------------------------------------------
    @torch.no_grad()
    def _rollout_body(self):
        env, buf, hp = self.env, self.buf, self.hp
        for t in range(self.n_steps):
            a, logp, v = self.policy.act(self._obs)
            if self.marl:
                # the reference's RLlib loop only ever sees agents that exist (rllib_envs/BaseAviary.py:425-430 build the
                # dicts over range(NUM_DRONES)); slots >= NUM_DRONES of a fixed-shape batch are padding, not samples
                n = env.num_drones_dev()
                buf["w"][t].copy_((self._slot[None, :] < n[:, None]).reshape(self.B))
            buf["obs"][t].copy_(self._obs)
            buf["act"][t].copy_(a)
            buf["logp"][t].copy_(logp)
            buf["val"][t].copy_(v)
            # SB3 clips to the action space only on the way into the env; the buffer keeps the unclipped sample
            obs, rew, _, _ = env.step(a.clamp(-1.0, 1.0).reshape(env.num_envs, env.max_drones, env.action_width).contiguous())
            done, boot = self._sample_flags()
            # time-limit truncation: bootstrap with the value of the terminal observation
            # (OnPolicyAlgorithm.collect_rollouts: rewards[idx] += gamma * predict_values(terminal_observation))
            tv = self.policy.value(self._flat_obs(env.terminal_obs))
            # where(): a stale terminal_obs row of an env that did not finish may hold anything, and NaN * 0 is NaN
            buf["rew"][t].copy_(rew.reshape(self.B) + hp["gamma"] * torch.where(boot > 0, tv, torch.zeros_like(tv)))
            buf["done"][t].copy_(done)
            self._obs.copy_(self._flat_obs(obs))      # episode statistics come from the simulator's accumulator (env.stats)
        last_v = self.policy.value(self._obs)
        adv, ret = gae(buf["rew"], buf["val"], buf["done"], last_v, hp["gamma"], hp["gae_lambda"])
        buf["adv"].copy_(adv)
        buf["ret"].copy_(ret)

    def collect(self):
        env = self.env
        if self._obs is None:
            self._obs = self._flat_obs(env.reset()).clone()
        if self.graph_rollout:
            if self._g_roll is None:
                side = torch.cuda.Stream(device=env.device)
                side.wait_stream(torch.cuda.current_stream(env.device))
                with torch.cuda.stream(side):
                    self._rollout_body()                   # warm-up outside the capture (allocator, cuBLAS handles)
                torch.cuda.current_stream(env.device).wait_stream(side)
                torch.cuda.synchronize(env.device)
                self._g_roll = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._g_roll, capture_error_mode="thread_local"):
                    self._rollout_body()
            else:
                self._g_roll.replay()
        else:
            self._rollout_body()
        self.num_timesteps += self.n_steps * env.num_envs
        return self.buf["adv"], self.buf["ret"]

    # ---- update (PPO.train) --------------------------------------------------------------------------------------
    def _minibatch_step(self, idx):
        hp, buf = self.hp, self.buf
        N = self.n_steps * self.B
        obs, act = buf["obs"].reshape(N, -1), buf["act"].reshape(N, -1)
        loss, pg, vf, ent, clipfrac = ppo_loss(self.policy, obs[idx], act[idx], buf["logp"].reshape(N)[idx], buf["adv"].reshape(N)[idx],
                                               buf["ret"].reshape(N)[idx], buf["w"].reshape(N)[idx], hp)
        self.opt.zero_grad(set_to_none=True)
        loss.backward()
        average_gradients(self._params)
        if hp["max_grad_norm"]:
            nn.utils.clip_grad_norm_(self._params, hp["max_grad_norm"])
        self.opt.step()
        with torch.no_grad():
            self._acc += torch.stack([pg.detach(), vf.detach(), ent.detach(), clipfrac.detach()])

    # ---- per-GPU learners: the same minibatch step cut in two around the gradient all-reduce ------------------------------
    def _bind_flat_grads(self):
        """Gradients of all parameters live in ONE flat buffer (p.grad are views of it): the all-reduce needs no packing
        and both halves of the step see fixed addresses, which is what lets them be captured separately."""
        n = sum(p.numel() for p in self._params)
        self._flat = torch.zeros((n,), dtype=torch.float32, device=self.env.device)
        o = 0
        for p in self._params:
            p.grad = self._flat[o:o + p.numel()].view_as(p)
            o += p.numel()

    def _mb_forward_backward(self, idx):
        hp, buf = self.hp, self.buf
        N = self.n_steps * self.B
        obs, act = buf["obs"].reshape(N, -1), buf["act"].reshape(N, -1)
        loss, pg, vf, ent, clipfrac = ppo_loss(self.policy, obs[idx], act[idx], buf["logp"].reshape(N)[idx], buf["adv"].reshape(N)[idx],
                                               buf["ret"].reshape(N)[idx], buf["w"].reshape(N)[idx], hp)
        self._flat.zero_()                 # backward accumulates into the views
        loss.backward()
        with torch.no_grad():
            self._acc += torch.stack([pg.detach(), vf.detach(), ent.detach(), clipfrac.detach()])

    def _mb_apply(self):
        import torch.distributed as dist
        with torch.no_grad():
            self._flat.mul_(1.0 / dist.get_world_size())
        if self.hp["max_grad_norm"]:
            nn.utils.clip_grad_norm_(self._params, self.hp["max_grad_norm"])
        self.opt.step()

    def _minibatch_step_multi(self, idx):
        import torch.distributed as dist
        self._mb_forward_backward(idx)
        dist.all_reduce(self._flat)
        self._mb_apply()

    def _capture_update_multi(self):
        """Two captures sharing the flat gradient buffer; the eager all-reduce between their replays is stream-ordered."""
        import torch.distributed as dist
        dev = self.env.device
        saved_p = [p.detach().clone() for p in self._params]
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                self._minibatch_step_multi(self._idx)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._g_upd, self._g_upd_b = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._g_upd, capture_error_mode="thread_local"):
            self._mb_forward_backward(self._idx)
        dist.all_reduce(self._flat)
        with torch.cuda.graph(self._g_upd_b, capture_error_mode="thread_local"):
            self._mb_apply()
        with torch.no_grad():                       # undo the warm-up and the capture passes
            for p, q in zip(self._params, saved_p):
                p.copy_(q)
            for st in self.opt.state.values():
                for v in st.values():
                    if torch.is_tensor(v):
                        v.zero_()
        torch.cuda.synchronize(dev)

    def _capture_update(self):
        """Whole-step capture (PyTorch CUDA-graph recipe): a few eager steps on a side stream, whose effect on the
        weights and the optimizer state is undone, then one captured step that every minibatch replays."""
        dev = self.env.device
        saved_p = [p.detach().clone() for p in self._params]
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):
                self._minibatch_step(self._idx)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._g_upd = torch.cuda.CUDAGraph()
        self.opt.zero_grad(set_to_none=True)
        # thread-local capture mode: the NCCL watchdog thread of a multi-GPU job may touch the CUDA API meanwhile
        with torch.cuda.graph(self._g_upd, capture_error_mode="thread_local"):
            self._minibatch_step(self._idx)
        with torch.no_grad():                       # undo the warm-up and the capture pass
            for p, q in zip(self._params, saved_p):
                p.copy_(q)
            for st in self.opt.state.values():
                for v in st.values():
                    if torch.is_tensor(v):
                        v.zero_()

    def update(self, adv=None, ret=None):
        hp, buf = self.hp, self.buf
        N = self.n_steps * self.B
        self._params = list(self.policy.parameters())
        dev = self.env.device
        graphs = self.cuda_graphs and N % self.batch_size == 0
        if self.multi and self._flat is None:
            self._bind_flat_grads()
        if graphs and self._g_upd is None:
            if self.multi:
                self._capture_update_multi()
            else:
                self._capture_update()
        self._acc.zero_()
        n_mb = 0
        import torch.distributed as dist
        for _ in range(hp["n_epochs"]):
            perm = torch.randperm(N, device=dev)
            for s in range(0, N, self.batch_size):
                if graphs and self.multi:
                    self._idx.copy_(perm[s:s + self.batch_size])
                    self._g_upd.replay()              # forward + backward into the flat gradient buffer
                    dist.all_reduce(self._flat)       # NCCL over NVLink, eager, stream-ordered between the two graphs
                    self._g_upd_b.replay()            # 1/world, clip, Adam
                elif graphs:
                    self._idx.copy_(perm[s:s + self.batch_size])
                    self._g_upd.replay()
                elif self.multi:
                    self._minibatch_step_multi(perm[s:s + self.batch_size])
                else:
                    self._minibatch_step(perm[s:s + self.batch_size])
                n_mb += 1
        acc = (self._acc / max(1, n_mb)).tolist()
        self.last = dict(pg=acc[0], vf=acc[1], ent=acc[2], clipfrac=acc[3])
        live = buf["w"].reshape(N) > 0
        ret_f, val_f = buf["ret"].reshape(N)[live], buf["val"].reshape(N)[live]
        self.last["explained_variance"] = float(1 - (ret_f - val_f).var() / (ret_f.var() + 1e-8))
        return self.last

    def learn(self, iterations, log=None):
        """`iterations` x (rollout + update).  Returns a list of per-iteration dicts (losses, fps, episode statistics
        from the simulator's on-device accumulator)."""
        out = []
        dev = self.env.device
        for it in range(iterations):
            self.env.stats(reset=True)
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            self.collect()
            torch.cuda.synchronize(dev)
            t1 = time.perf_counter()
            info = dict(self.update())
            torch.cuda.synchronize(dev)
            t2 = time.perf_counter()
            st = self.env.stats()
            steps = self.n_steps * self.env.num_envs
            info.update(iteration=it, timesteps=self.num_timesteps, rollout_s=t1 - t0, update_s=t2 - t1,
                        rollout_env_steps_per_s=steps / (t1 - t0), env_steps_per_s=steps / (t2 - t0),
                        episodes=float(st[0]), ep_rew_mean=float(st[1] / st[0]) if st[0] else float("nan"),
                        ep_len_mean=float(st[2] / st[0]) if st[0] else float("nan"))
            out.append(info)
            if log:
                log(info)
        return out
