"""Does step time scale with the number of resident CTAs per SM?  Pads the dynamic shared memory (CHS_SMEM_PAD)
so that 4, 3, 2, 1 CTAs of the (G=32, 256 threads) shape fit per SM and times 65536 envs."""
import os
import subprocess
import sys

code = r'''
import os, sys, torch, numpy as np
sys.path.insert(0, os.getcwd())
import rl_cattle_herding_b200 as pkg
E, D, M = 65536, 3, 20
env = pkg.CattleHerdingBatchedEnv(E, num_drones=D, num_cattle=M, ctrl_freq=30, fixed_num_drones=D,
                                  spawn_table=pkg.spawn.generate(100, M, seed=0), device=0, seed=1)
env.set_launch(int(os.environ["G"]), int(os.environ["TH"]))
pkg._lib.check(env.lib.chs_set_register_variant(env._h, int(os.environ["VAR"])))
env.reset()
acts = [torch.rand((E, D, 4), device="cuda") * 2 - 1 for _ in range(4)]
for i in range(4): env.step(acts[i])
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
for i, (a, b) in enumerate(ev):
    a.record(); env.step(acts[i % 4]); b.record()
torch.cuda.synchronize()
t = [a.elapsed_time(b) for a, b in ev]
print(os.environ.get("CHS_SMEM_PAD", "0"), os.environ["G"], os.environ["TH"], os.environ["VAR"], env.dims["smem_bytes"],
      "even %.1f us odd %.1f us" % (np.mean(t[0::2]) * 1e3, np.mean(t[1::2]) * 1e3))
'''
for g, th, var, pads in ((32, 256, 2, (0, 12000, 30000, 70000)), (32, 96, 1, (0, 12000, 30000, 70000)), (16, 128, 2, (0, 5000, 14000, 32000, 52000))):
    for pad in pads:
        env = dict(os.environ, CHS_SMEM_PAD=str(pad), G=str(g), TH=str(th), VAR=str(var))
        subprocess.run([sys.executable, "-c", code], env=env)
