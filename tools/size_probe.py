"""Fused-kernel step time vs batch size at a fixed CTA shape (GPU box): separates the latency of one CTA from what
co-resident CTAs add.  usage: python tools/size_probe.py G threads E [E ...]"""
import os
import sys
import torch

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import rl_cattle_herding_b200 as pkg  # noqa: E402
from launch_sweep import time_shape  # noqa: E402

G, TH = int(sys.argv[1]), int(sys.argv[2])
flush = None if os.environ.get("NO_FLUSH") else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for E in [int(x) for x in sys.argv[3:]]:
    env = pkg.CattleHerdingBatchedEnv(E, num_drones=3, num_cattle=20, ctrl_freq=30, fixed_num_drones=3,
                                      spawn_table=pkg.spawn.generate(100, 20, seed=0), device=0, seed=1,
                                      auto_reset=os.environ.get("NO_RESET") is None)
    env.reset()
    env.set_step_mode("fused")
    env.set_launch(G, TH)
    acts = [torch.rand((E, 3, 4), device="cuda") * 2 - 1 for _ in range(4)]
    for i in range(30):
        env.step(acts[i % 4])
    ms = time_shape(env, acts, flush, 24)
    print(f"E={E:6d} G={G} threads={TH} CTAs={(E + G - 1) // G:5d}: {ms * 1e3:7.2f} us")
    env.close()
