"""CPU arm of bench.py: times the reference's OWN code (CattleAviary, Physics.DYN) on the host cores.
TEST / BASELINE INFRASTRUCTURE ONLY -- nothing here is on the product path.

Two structures, both asked for by BASELINE.json's north_star ("single-process and SubprocVecEnv"):

  * `time_single`: one process, one `CattleAviary` imported unmodified through `oracle/ref_harness/ref_env.py`,
    uniform random VEL actions, `reset()` whenever an episode ends -- the manual loop of
    simulator/CTDECattleHerder.py:198-220.
  * `time_pool`: P worker processes, one env each, driven the way stable_baselines3's SubprocVecEnv drives them
    (simulator/CTDECattleHerder.py:91-97: `make_vec_env(..., n_envs=24, vec_env_cls=SubprocVecEnv)`): per control step
    the parent sends every worker its action over a pipe, the worker steps, resets itself when the episode ended
    (keeping the terminal observation) and sends (obs, reward, done, info) back; the parent stacks the P results.
    stable_baselines3 itself is not installed on either box, so the protocol is restated here.

Falls back to the float64 port (`cattle_oracle.OracleEnv`, kind "port") when neither /root/reference nor the staged
copy `oracle/_ref` (oracle/make_ref.py) is present.
"""
import multiprocessing as mp
import os
import sys
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
for _p in (_HERE, os.path.join(_HERE, "ref_harness")):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def kind():
    import ref_env
    return "reference" if ref_env.reference_available() else "port"


def host_info():
    model = "unknown"
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    model = line.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    return {"cpu_model": model, "cores_visible": cores}


def third_party():
    import ref_env
    return ref_env.third_party_kinds() if ref_env.reference_available() else {}


class _Env:
    """One env behind a uniform step/reset surface: the reference if it can be imported, else the port."""

    def __init__(self, n, m, ctrl_freq, table, seed):
        import ref_env
        self.n, self.m = n, m
        self.rng = np.random.default_rng(seed)
        if ref_env.reference_available():
            self.ref = ref_env
            self.env = ref_env.make_reference_env(n, m, ctrl_freq=ctrl_freq, spawn_table=table, seed=seed)
            self.port = None
        else:
            import cattle_oracle as co
            self.ref = None
            self.port = co.OracleEnv(m, ctrl_freq=ctrl_freq, spawn_table=table, start_level=7)

    def reset(self):
        if self.ref:
            return self.ref.quiet(self.env.reset)[0]
        return self.port.reset(self.n, self.rng.random((self.m, 2)))

    def step(self, a):
        if self.ref:
            obs, rew, term, trunc, info = self.ref.quiet(self.env.step, a)
        else:
            obs, rew, term, trunc, info = self.port.step(a)
        return obs, float(rew), bool(term), bool(trunc), info


def _table(m):
    import ref_env
    return ref_env.synthetic_spawn_table(16 if m <= 32 else 4, m, seed=0)


def time_single(n, m, ctrl_freq=30, budget_s=10.0, seed=1):
    """Returns (env-steps/s, description)."""
    env = _Env(n, m, ctrl_freq, _table(m), seed)
    env.reset()
    rng = np.random.default_rng(seed)
    steps = episodes = 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < budget_s:
        _, _, term, trunc, _ = env.step(rng.uniform(-1, 1, (n, 4)).astype(np.float32))
        steps += 1
        if term or trunc:
            env.reset()
            episodes += 1
    dt = time.perf_counter() - t0
    return steps / dt, f"1 process x 1 env, {steps} control steps, {episodes} episodes in {dt:.1f} s"


def _worker(conn, n, m, ctrl_freq, seed):
    env = _Env(n, m, ctrl_freq, _table(m), seed)
    obs = env.reset()
    conn.send(("ready", np.asarray(obs, dtype=np.float32)))
    while True:
        cmd, data = conn.recv()
        if cmd == "step":
            obs, rew, term, trunc, info = env.step(data)
            done = term or trunc
            info = dict(info) if isinstance(info, dict) else {}
            info["TimeLimit.truncated"] = trunc and not term
            if done:
                info["terminal_observation"] = np.asarray(obs, dtype=np.float32)
                obs = env.reset()
            conn.send((np.asarray(obs, dtype=np.float32), rew, done, info))
        elif cmd == "close":
            conn.close()
            return


def time_pool(n, m, ctrl_freq=30, budget_s=15.0, procs=None, seed=100):
    """Returns (aggregate env-steps/s, P, description): P processes, one pipe round trip per env per step."""
    procs = procs or host_info()["cores_visible"]
    ctx = mp.get_context("fork")
    conns, ps = [], []
    for i in range(procs):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_worker, args=(b, n, m, ctrl_freq, seed + i), daemon=True)
        p.start()
        b.close()
        conns.append(a)
        ps.append(p)
    for c in conns:
        c.recv()
    rng = np.random.default_rng(seed)
    steps = 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < budget_s:
        acts = rng.uniform(-1, 1, (procs, n, 4)).astype(np.float32)
        for c, a in zip(conns, acts):          # step_async
            c.send(("step", a))
        res = [c.recv() for c in conns]        # step_wait
        np.stack([r[0] for r in res])
        np.array([r[1] for r in res], dtype=np.float32)
        steps += 1
    dt = time.perf_counter() - t0
    for c in conns:
        c.send(("close", None))
    for p in ps:
        p.join(timeout=5)
    return procs * steps / dt, procs, (f"{procs} processes x 1 env, pipe round trip per step (SubprocVecEnv protocol), "
                                       f"{steps} vector steps = {procs * steps} env-steps in {dt:.1f} s")
