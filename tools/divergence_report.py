"""Free-running divergence of the fp32 CUDA simulator from the float64 oracle over a 1000-step rollout
(BASELINE.json north_star: "divergence bounds reported over a 1000-step rollout in fp32").  GPU box.

Both sides start every episode from the same state (the device's post-reset state is injected into the oracle:
RNG streams are the device's) and then run freely on identical actions until the episode ends on the oracle
side; at that point both are reset and re-synchronised.  Three action laws: zero (hover), gentle
(|a3| <= 0.05) and uniform.   python tools/divergence_report.py [steps] [envs] > profiles/r1_divergence_fp32.txt
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import cattle_oracle as co  # noqa: E402
import rl_cattle_herding_b200 as pkg  # noqa: E402

STEPS = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
E = int(sys.argv[2]) if len(sys.argv) > 2 else 8
N, M = 3, 20
table = pkg.spawn.generate(100, M, seed=0)


def pad(a):
    a = np.asarray(a)
    if a.ndim == 2 and a.shape[0] < 12 and a.shape[1] in (3, 4):
        out = np.zeros((12, a.shape[1])); out[: a.shape[0]] = a
        return out
    return a


def sync(env, oracles, idx):
    st = env.get_state()
    for e in idx:
        oracles[e].set_state({k: pad(st[k][e]) for k in co.STATE_FIELDS})


for law in ("zero", "gentle", "uniform"):
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, ctrl_freq=30, fixed_num_drones=N, auto_reset=False,
                                      spawn_table=table, device=0, seed=3)
    env.reset()
    oracles = [co.OracleEnv(M, ctrl_freq=30, spawn_table=table) for _ in range(E)]
    sync(env, oracles, range(E))
    rng = np.random.default_rng(7)
    age = np.zeros(E, dtype=int)
    by_age = {}            # steps since resync -> list of max |pos| divergence
    worst = dict(pos=0.0, vel=0.0, rpy=0.0, obs=0.0, reward=0.0, cattle_vel=0.0)
    flag_mismatch, episodes, first_mismatch = 0, 0, None
    for t in range(STEPS):
        a = rng.uniform(-1, 1, (E, N, 4)).astype(np.float32)
        if law == "zero":
            a[:] = 0
        elif law == "gentle":
            a[..., 3] = np.abs(a[..., 3]) * 0.05
        obs, rew, term, trunc = env.step(torch.from_numpy(a).cuda())
        obs, rew = obs.cpu().numpy(), rew.cpu().numpy()
        done_dev = (term | trunc).bool().cpu().numpy()
        st = env.get_state()
        age += 1
        resync = []
        for e in range(E):
            o = oracles[e]
            oobs, orew, oterm, otrunc, _ = o.step(a[e])
            d = dict(pos=np.abs(st["pos"][e] - o.pos[:N]).max(), vel=np.abs(st["vel"][e] - o.vel[:N]).max(),
                     rpy=np.abs(st["rpy"][e] - o.rpy[:N]).max(), obs=np.abs(obs[e] - oobs).max(),
                     reward=abs(float(rew[e]) - orew), cattle_vel=np.abs(st["cattle_vel"][e] - o.cattle_vel).max())
            for k, v in d.items():
                if np.isfinite(v):
                    worst[k] = max(worst[k], float(v))
            by_age.setdefault(int(age[e]), []).append(float(d["pos"]))
            odone = oterm or otrunc
            if bool(done_dev[e]) != bool(odone):
                flag_mismatch += 1
                first_mismatch = first_mismatch if first_mismatch is not None else (t, e, int(age[e]))
            if odone or done_dev[e]:
                episodes += 1
                resync.append(e)
        if resync:
            mask = torch.zeros(E, dtype=torch.uint8)
            mask[resync] = 1
            env.reset(mask=mask.cuda())
            sync(env, oracles, resync)
            age[resync] = 0
    print(f"== action law: {law}; {E} envs x {STEPS} steps, {N} drones / {M} cattle, fp32 build vs float64 oracle ==")
    print(f"episodes finished: {episodes}; env-steps with a differing done flag: {flag_mismatch} of {E * STEPS}"
          f" (first at step/env/age {first_mismatch})")
    print("max |cuda - oracle| over the whole rollout: " + ", ".join(f"{k}={v:.2e}" for k, v in worst.items()))
    ages = sorted(by_age)
    pick = [x for x in (1, 2, 4, 8, 12, 16, 24, 32, 64, 128, 256, 512, 1000) if x in by_age]
    print("max position divergence by steps since resync: " + ", ".join(f"{x}:{max(by_age[x]):.1e}" for x in pick))
    print(f"longest free run: {ages[-1]} steps")
    env.close()
