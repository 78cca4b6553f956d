"""Workload for ncu captures (GPU box): steps one env batch into its steady state, then a few more steps for ncu to capture.
    ncu --set full --clock-control none --import-source on -k regex:step_kernel --launch-skip 60 --launch-count 2 \
        -o gpurun_out/prof python tools/ncu_target.py E N M mode G T
"""
import os
import sys

import torch

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import rl_cattle_herding_b200 as pkg  # noqa: E402

E, N, M = (int(x) for x in sys.argv[1:4])
mode = sys.argv[4] if len(sys.argv) > 4 else "auto"
G = int(sys.argv[5]) if len(sys.argv) > 5 else 0
T = int(sys.argv[6]) if len(sys.argv) > 6 else 0
env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, fixed_num_drones=N, spawn_table=pkg.spawn.generate(100 if M <= 32 else 8, M, seed=0),
                                  device=0, seed=1, step_mode=mode, envs_per_block=G, threads_per_block=T)
env.reset()
acts = [torch.rand((E, N, 4), device="cuda") * 2 - 1 for _ in range(4)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if E * 1400 < 120e6 else None
for i in range(int(os.environ.get("NCU_STEPS", "66"))):
    if flush is not None:
        flush.zero_()
    env.step(acts[i % 4])
torch.cuda.synchronize()
print("dims", env.dims)
