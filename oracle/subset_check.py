"""Subset parity check of a running batched simulator against the CPU oracle.  TEST INFRASTRUCTURE ONLY.

Used by `tests/test_cuda_shapes.py` and by the untimed `parity_check` stamp of `bench.py`: before a step the
state of a strided subset of envs is pulled off the device (`chs_get_state`), after the step every one of them is
advanced by `cattle_oracle.OracleEnv` (float64 restatement of sb3_envs/BaseAviary.py:335-465, pinned to the
reference's recorded outputs by tests/test_oracle_golden.py) from that state with the same actions, and the
device's observation / reward / flags -- and, for envs the device auto-reset, the first observation of the new
episode (BaseAviary.py:280-331 fed with the device's own Philox draws) -- have to agree:

  * terminated / truncated: bit exact
  * observation: |err| <= obs_rel * max(1, |ref|_inf) + 1e-6   (fp32 build: obs_rel = 1e-5, tests/test_cuda_parity.py)
  * reward:      |err| <= rew_tol * max(1, |ref|)              (fp32 build: 5e-5, same file)

Neighbour columns of rows whose ranking hinges on a distance gap below fp32 resolution are compared up to the
choice among equidistant drones (tests/_golden.py: near_tie_rows has the reasoning).
"""
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
if _HERE not in sys.path:
    sys.path.insert(0, _HERE)

import cattle_oracle as co  # noqa: E402

NMAX = co.NMAX


def subset_ids(num_envs, n_blocks=32, block=8):
    """`n_blocks` runs of `block` consecutive envs spread over the batch, first and last run included."""
    block = min(block, num_envs)
    n_blocks = max(1, min(n_blocks, num_envs // block))
    if n_blocks == 1:
        starts = [0]
    else:
        last = num_envs - block
        starts = sorted({int(round(i * last / (n_blocks - 1))) for i in range(n_blocks)})
    ids = []
    for s in starts:
        ids.extend(range(s, s + block))
    return sorted(set(ids))


def _runs(ids):
    runs, start, prev = [], ids[0], ids[0]
    for e in ids[1:]:
        if e != prev + 1:
            runs.append((start, prev - start + 1))
            start = e
        prev = e
    runs.append((start, prev - start + 1))
    return runs


def _pad(a):
    a = np.asarray(a)
    if a.ndim == 2 and a.shape[1] in (3, 4) and a.shape[0] < NMAX:
        out = np.zeros((NMAX, a.shape[1]))
        out[: a.shape[0]] = a
        return out
    return a


def _near_tie_rows(pos, n, tol=2e-6):
    pos = np.asarray(pos, dtype=np.float64)[:n, :2]
    out = np.zeros((pos.shape[0],), dtype=bool)
    for i in range(n):
        d = np.sort(np.delete(np.sqrt(np.sum((pos - pos[i]) ** 2, axis=1)), i))
        gaps = np.diff(d[:3])
        out[i] = bool(np.any(gaps < tol * max(1.0, np.abs(pos).max())))
    return out


def _mask_ties(obs, ref, pos, n, tol=2e-5):
    obs = obs.copy()
    tie = _near_tie_rows(pos, n)
    xy = np.asarray(pos, dtype=np.float64)[:n, :2]
    for i in np.nonzero(tie)[0]:
        rel = np.delete(xy - xy[i], i, axis=0)
        for c in (10, 12):
            ours, theirs = obs[i, c:c + 2].astype(np.float64), ref[i, c:c + 2].astype(np.float64)
            if rel.size and np.min(np.linalg.norm(rel - ours, axis=1)) <= tol and \
                    abs(np.linalg.norm(ours) - np.linalg.norm(theirs)) <= tol:
                obs[i, c:c + 2] = ref[i, c:c + 2]
    return obs


class SubsetChecker:
    """`env`: a CattleHerdingBatchedEnv (CTDE packing, obs_rows 12 or max_drones).  `ids`: local env indices."""

    def __init__(self, env, table, ids, seed=0, env_id_offset=0, start_level=7, fixed_num_drones=0,
                 obs_rel=1e-5, rew_tol=5e-5):
        self.env, self.table, self.ids = env, np.asarray(table, dtype=np.float64), list(ids)
        self.seed, self.off, self.level, self.fixed_n = seed, env_id_offset, start_level, fixed_num_drones
        self.obs_rel, self.rew_tol = obs_rel, rew_tol
        self.state = None
        self.n_checked = self.n_done = 0
        self.max_obs_err = self.max_rew_err = self.max_reset_obs_err = 0.0
        self.flags_equal = True
        self.failures = []

    def before(self):
        parts = [self.env.get_state(s, c) for s, c in _runs(self.ids)]
        self.state = {k: np.concatenate([p[k] for p in parts], axis=0) for k in parts[0]}

    def after(self, actions, obs, reward, terminated, truncated, terminal_obs=None):
        """All arguments are host numpy arrays GATHERED for the subset: row i belongs to env self.ids[i]
        (`gather` below does that on the device, so that a 262 144-env observation tensor never crosses PCIe)."""
        from rl_cattle_herding_b200 import philox as ph
        env, st = self.env, self.state
        R = env.obs_rows
        for i, e in enumerate(self.ids):
            o = co.OracleEnv(env.num_cattle, ctrl_freq=env.ctrl_freq, pyb_freq=env.pyb_freq, spawn_table=self.table,
                             start_level=self.level)
            o.set_state({k: _pad(st[k][i]) for k in co.STATE_FIELDS})
            n = int(st["n_drones"][i])
            oobs, orew, oterm, otrunc, _ = o.step(np.asarray(actions[i], dtype=np.float32)[:max(n, 1)])
            t, u = bool(np.asarray(terminated[i]).item()), bool(np.asarray(truncated[i]).item())
            if (t, u) != (bool(oterm), bool(otrunc)):
                self.flags_equal = False
                self.failures.append(("flags", e, (t, u), (bool(oterm), bool(otrunc))))
            done = t or u
            mine = np.asarray(terminal_obs[i] if (done and env.auto_reset and terminal_obs is not None) else obs[i])
            if done and env.auto_reset and terminal_obs is None:
                mine = None
            if mine is not None:
                ref = np.asarray(oobs)[:R]
                mine = _mask_ties(mine, ref, o.pos, n)
                err = float(np.max(np.abs(mine.astype(np.float64) - ref)))
                tol = self.obs_rel * max(1.0, float(np.max(np.abs(ref)))) + 1e-6
                self.max_obs_err = max(self.max_obs_err, err)
                if not err <= tol:
                    self.failures.append(("obs", e, err, tol))
            if np.isfinite(orew):
                rerr = abs(float(np.asarray(reward[i]).item()) - orew)
                self.max_rew_err = max(self.max_rew_err, rerr)
                if not rerr <= self.rew_tol * max(1.0, abs(orew)):
                    self.failures.append(("reward", e, rerr, orew))
            if done and env.auto_reset:
                # the device rebuilt this env inside the step: oracle reset with the device's own draws
                gid, epi = self.off + e, int(st["episode"][i]) + 1
                nn = self.fixed_n if self.fixed_n > 0 else ph.num_drones(self.seed, gid, epi, o.MIN_NUM_DRONES,
                                                                         o.MAX_NUM_DRONES, env.max_drones)
                robs = np.asarray(o.reset(nn, ph.cow_uniforms(self.seed, gid, epi, env.num_cattle)))[:R]
                err = float(np.max(np.abs(np.asarray(obs[i]).astype(np.float64) - robs)))
                self.max_reset_obs_err = max(self.max_reset_obs_err, err)
                if not err <= 4e-5:
                    self.failures.append(("reset_obs", e, err))
                self.n_done += 1
            self.n_checked += 1

    def gather(self, *tensors):
        """Rows of the subset of each CUDA tensor, as numpy arrays (None passes through)."""
        import torch
        idx = torch.as_tensor(self.ids, dtype=torch.long, device=self.env.device)
        return [None if t is None else t.index_select(0, idx).cpu().numpy() for t in tensors]

    def step(self, actions, want_terminal_obs=True):
        """before() + env.step(actions) + after() in one call; returns what env.step returned."""
        self.before()
        out = self.env.step(actions, want_terminal_obs=want_terminal_obs)
        tobs = self.env.terminal_obs if (want_terminal_obs and self.env.auto_reset) else None
        self.after(*self.gather(actions, out[0], out[1], out[2], out[3], tobs))
        return out

    def summary(self):
        return {"envs": len(self.ids), "env_steps": self.n_checked, "resets_checked": self.n_done,
                "max_obs_err": self.max_obs_err, "max_reward_err": self.max_rew_err,
                "max_reset_obs_err": self.max_reset_obs_err, "flags_equal": bool(self.flags_equal),
                "ok": not self.failures}
