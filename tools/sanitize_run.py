"""Small workload for compute-sanitizer (memcheck / racecheck / synccheck / initcheck) on the GPU box:
    compute-sanitizer --tool racecheck python tools/sanitize_run.py [fused|split|static|host|reset|c3|marl ...]
512 envs, auto-reset on, enough steps for envs to finish and be rebuilt inside the kernels (bar.arrive / named barriers,
bulk-async stores, fence.proxy.async, in-kernel reset: csrc/chs_kernel.cuh phases 0-5)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import rl_cattle_herding_b200 as pkg  # noqa: E402

what = sys.argv[1:] or ["fused", "split", "static", "host", "reset", "c3", "marl"]
E = int(os.environ.get("SAN_ENVS", "512"))
STEPS = int(os.environ.get("SAN_STEPS", "24"))


def make(N=3, M=20, **kw):
    env = pkg.CattleHerdingBatchedEnv(E if M <= 32 else E // 4, num_drones=N, num_cattle=M, fixed_num_drones=kw.pop("fixed", N),
                                      spawn_table=pkg.spawn.generate(8, M, seed=0), device=0, seed=3, **kw)
    env.reset()
    return env


def drive(env, call, steps=STEPS):
    gen = torch.Generator("cuda").manual_seed(0)
    done = 0
    for _ in range(steps):
        a = torch.rand((env.num_envs, env.max_drones, 4), device="cuda", generator=gen) * 2 - 1
        out = call(a)
        done += int((out[2].reshape(env.num_envs, -1)[:, -1] | out[3].reshape(env.num_envs, -1)[:, -1]).sum())
    torch.cuda.synchronize()
    return done


for w in what:
    if w == "fused":
        for g, th in ((8, 128), (16, 256), (4, 64)):
            env = make(step_mode="fused", envs_per_block=g, threads_per_block=th)
            print(f"fused G={g} T={th}: finished {drive(env, env.step)}", flush=True)
            env.close()
    elif w == "split":
        env = make(step_mode="split", envs_per_block=32, threads_per_block=256)
        print(f"split: finished {drive(env, env.step)}", flush=True)
        env.close()
    elif w == "static":
        env = make(step_mode="split")
        print(f"static scratch: finished {drive(env, env.step_static_scratch)}", flush=True)
        env.close()
    elif w == "host":
        env = make()
        env.reset_host()
        rng = np.random.default_rng(0)
        n = 0
        for _ in range(STEPS):
            n += len(env.step_host(rng.uniform(-1, 1, (env.num_envs, 3, 4)).astype(np.float32))[4])
        print(f"host path (direct to pinned): finished {n}", flush=True)
        env.close()
    elif w == "reset":
        env = make(auto_reset=False, num_drones=12, fixed=0, N=12, M=16)
        for i in range(4):
            m = (torch.arange(env.num_envs, device="cuda") % (i + 2) == 0).to(torch.uint8)
            env.reset(mask=m)
        torch.cuda.synchronize()
        print("masked reset ok", flush=True)
        env.close()
    elif w == "c3":
        env = make(8, 100, step_mode="fused")
        print(f"c3 fused: finished {drive(env, env.step, 12)}", flush=True)
        env.close()
    elif w == "marl":
        env = make(4, 20, packing="marl", start_level=0)
        print(f"marl: finished {drive(env, env.step)}", flush=True)
        env.close()
