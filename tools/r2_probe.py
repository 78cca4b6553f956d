"""Round-2 probe (GPU box): step time of the shapes bench.py reports, several ways, plus the host path.
    python tools/r2_probe.py [what ...]      what in {head, rot, big, c3, host}; default all
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import rl_cattle_herding_b200 as pkg  # noqa: E402

what = set(sys.argv[1:]) or {"head", "rot", "big", "c3", "host"}
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def make(E, N=3, M=20, **kw):
    table = pkg.spawn.generate(100 if M <= 32 else 8, M, seed=0)
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, fixed_num_drones=N, spawn_table=table, device=0, seed=0, **kw)
    env.reset()
    return env


def isolated(env, acts, n, fl=True):
    st = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    en = [torch.cuda.Event(enable_timing=True) for _ in range(n)]
    for i in range(n):
        if fl:
            flush.zero_()
        st[i].record()
        env.step(acts[i % len(acts)])
        en[i].record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in zip(st, en))
    return 1e3 * float(np.mean(ts[: max(1, 3 * n // 4)]))


def acts_for(E, N, k=4):
    return [torch.rand((E, N, 4), device=dev) * 2 - 1 for _ in range(k)]


if "head" in what:
    for g, th in ((4, 64), (8, 96), (8, 128), (16, 128), (16, 256)):
        env = make(4096, envs_per_block=g, threads_per_block=th, step_mode="fused")
        a = acts_for(4096, 3, 8)
        isolated(env, a, 30)
        print(f"head 4096 fused G={g} T={th}: isolated+flush {isolated(env, a, 100):.2f} us, isolated no flush {isolated(env, a, 100, False):.2f} us", flush=True)
        env.close()

if "rot" in what:
    # back-to-back launches over R independent 4096-env batches (working set R x 17 MB > L2), one event pair
    for R in (1, 12):
        envs = [make(4096, envs_per_block=8, threads_per_block=128, step_mode="fused") for _ in range(R)]
        a = acts_for(4096, 3, 8)
        for i in range(40 * R):
            envs[i % R].step(a[i % 8])
        torch.cuda.synchronize()
        K = 600
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for i in range(K):
            envs[i % R].step(a[i % 8])
        e.record()
        torch.cuda.synchronize()
        print(f"rot 4096 x {R} batches back-to-back: {1e3 * s.elapsed_time(e) / K:.2f} us per step", flush=True)
        # same through a CUDA graph of 48 steps
        g = torch.cuda.CUDAGraph()
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for i in range(2 * R):
                envs[i % R].step(a[i % 8])
            with torch.cuda.graph(g, stream=st):
                for i in range(48):
                    envs[i % R].step(a[i % 8])
        torch.cuda.synchronize()
        g.replay()
        torch.cuda.synchronize()
        s.record()
        for _ in range(10):
            g.replay()
        e.record()
        torch.cuda.synchronize()
        print(f"rot 4096 x {R} batches, graph of 48 steps: {1e3 * s.elapsed_time(e) / 480:.2f} us per step", flush=True)
        for x in envs:
            x.close()

if "big" in what:
    for E in (65536, 262144):
        for pp in (True, False):
            for g, th in ((32, 256), (16, 256), (32, 128)):
                env = make(E, envs_per_block=g, threads_per_block=th, step_mode="split", persistent_padding=pp)
                a = acts_for(E, 3, 2)
                isolated(env, a, 30, False)
                print(f"big {E} split G={g} T={th} persistent_padding={pp}: {isolated(env, a, 40, False):.1f} us", flush=True)
                env.close()

if "c3" in what:
    for mode in ("fused", "split"):
        for g, th in ((2, 128), (4, 128), (4, 256), (8, 256)):
            env = make(1024, 8, 100, envs_per_block=g, threads_per_block=th, step_mode=mode)
            a = acts_for(1024, 8, 8)
            isolated(env, a, 30)
            print(f"c3 8d100c 1024 {mode} G={g} T={th}: {isolated(env, a, 60):.2f} us", flush=True)
            env.close()

if "host" in what:
    for nodirect in ("", "1"):
        if nodirect:
            os.environ["CHS_NO_DIRECT"] = "1"
        env = make(4096)
        ha = [x.cpu().numpy() for x in acts_for(4096, 3, 4)]
        env.reset_host()
        for i in range(10):
            env.step_host(ha[i % 4])
        t0 = time.perf_counter()
        K = 300
        for i in range(K):
            env.step_host(ha[i % 4])
        dt = time.perf_counter() - t0
        buf = env.host_actions
        t0 = time.perf_counter()
        for i in range(K):
            env.step_host()
        dt2 = time.perf_counter() - t0
        print(f"host 4096 {'staged' if nodirect else 'direct'}: {1e6 * dt / K:.1f} us per step ({4096 * K / dt:.3e} env-steps/s); "
              f"actions in place {1e6 * dt2 / K:.1f} us ({4096 * K / dt2:.3e})", flush=True)
        env.close()
