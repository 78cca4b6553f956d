"""What one step_host (numpy in / numpy out) spends where: CUPTI activity records of kernels and copies (GPU box)."""
import os
import sys
import time
import numpy as np
import torch
from torch.profiler import profile, ProfilerActivity

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import rl_cattle_herding_b200 as pkg  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
env = pkg.CattleHerdingBatchedEnv(E, num_drones=3, num_cattle=20, ctrl_freq=30, fixed_num_drones=3,
                                  spawn_table=pkg.spawn.generate(100, 20, seed=0), device=0, seed=1)
env.reset_host()
acts = [np.random.default_rng(i).uniform(-1, 1, (E, 3, 4)).astype(np.float32) for i in range(4)]
for i in range(50):
    env.step_host(acts[i % 4])
t0 = time.perf_counter()
for i in range(200):
    env.step_host(acts[i % 4])
wall = (time.perf_counter() - t0) / 200
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(40):
        env.step_host(acts[i % 4])
agg = {}
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        a = agg.setdefault(e.name[:70], [0, 0.0])
        a[0] += 1; a[1] += e.time_range.elapsed_us()
print(f"E={E}: {wall * 1e6:.1f} us per step_host (unprofiled wall clock); GPU-side activity per step:")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {t / 40:8.1f} us/step  ({n / 40:.1f} x {t / n:6.1f} us)  {k}")
