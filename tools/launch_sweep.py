"""Launch-shape sweep of the fused step kernel (GPU box): ms per step for (envs_per_block, threads, register variant).
usage: python tools/launch_sweep.py [E ...]   -- prints one line per shape, best first."""
import sys
import os
import numpy as np
import torch

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import rl_cattle_herding_b200 as pkg  # noqa: E402


def time_shape(env, acts, flush, reps):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i in range(3):
        env.step(acts[i % len(acts)])
    for i, (a, b) in enumerate(ev):
        if flush is not None:
            flush.zero_()
        a.record(); env.step(acts[i % len(acts)]); b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return float(np.mean(t[: max(1, len(t) * 3 // 4)]))


def main():
    sizes = [int(x) for x in sys.argv[1:]] or [4096, 65536]
    D, M = int(os.environ.get("CHS_D", 3)), int(os.environ.get("CHS_M", 20))
    table = pkg.spawn.generate(100, M, seed=0)
    for E in sizes:
        env = pkg.CattleHerdingBatchedEnv(E, num_drones=D, num_cattle=M, ctrl_freq=30, fixed_num_drones=D, spawn_table=table,
                                          device=0, seed=1)
        env.reset()
        acts = [torch.rand((E, D, 4), device="cuda") * 2 - 1 for _ in range(4)]
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if E * 4230 < (200 << 20) else None
        res = []
        for mode in ("fused", "split"):
            env.set_step_mode(mode)
            for g in (4, 8, 16, 32):
                for th in (64, 96, 128, 256, 512):
                    for var in (1, 2):
                        if var == 2 and th > 256:
                            continue
                        try:
                            env.set_launch(g, th)
                            pkg._lib.check(env.lib.chs_set_register_variant(env._h, var))
                            ms = time_shape(env, acts, flush, 16 if E <= 65536 else 8)
                            res.append((ms, g, th, var, mode))
                        except RuntimeError as ex:
                            print("skip", g, th, var, str(ex)[:60])
        res.sort()
        print(f"E={E} D={D} M={M}")
        for ms, g, th, var, sm in res[:12]:
            print(f"  {ms * 1e3:9.1f} us  G={g:2d} threads={th:3d} variant={var} {sm}")
        best_fused = min(r for r in res if r[4] == "fused")
        print("  best fused:", round(best_fused[0] * 1e3, 1), best_fused[1:])
        print("  worst:", [(round(r[0] * 1e3, 1), r[1], r[2], r[3]) for r in res[-3:]])
        env.close()


if __name__ == "__main__":
    main()
