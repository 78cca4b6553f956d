import os, sys, torch
sys.path.insert(0, os.getcwd())
import rl_cattle_herding_b200 as pkg
E, D, M = 262144, 3, 20
env = pkg.CattleHerdingBatchedEnv(E, num_drones=D, num_cattle=M, ctrl_freq=30, fixed_num_drones=D, spawn_table=pkg.spawn.generate(100, M, seed=0), device=0, seed=1, step_mode="split")
env.set_launch(32, 256); pkg._lib.check(env.lib.chs_set_register_variant(env._h, 2))
env.reset()
acts = [torch.rand((E, D, 4), device="cuda") * 2 - 1 for _ in range(4)]
for i in range(int(os.environ.get("NSTEPS", 12))): env.step(acts[i % 4])
torch.cuda.synchronize()
