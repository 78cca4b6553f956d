import ctypes as C, os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
import rl_cattle_herding_b200 as pkg
E, G, TH, N, M = 4096, 8, 128, 3, 20
env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, fixed_num_drones=N, spawn_table=pkg.spawn.generate(100, M), device=0, envs_per_block=G, threads_per_block=TH)
env.reset()
clk = torch.zeros(((E + G - 1) // G, 32), dtype=torch.int64, device="cuda")
acts = [torch.rand((E, N, 4), device="cuda") * 2 - 1 for _ in range(8)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(tag, n=40):
    ts = []
    for i in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); env.step(acts[i % 8]); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = sorted(ts)[: n * 3 // 4]
    print(tag, "%.2f us" % np.mean(ts))
for i in range(20): env.step(acts[i % 8])
run("stamps off")
pkg._lib.check(env.lib.chs_set_phase_clock(env._h, C.c_void_p(clk.data_ptr())))
run("stamps on ")
pkg._lib.check(env.lib.chs_set_phase_clock(env._h, C.c_void_p(0)))
run("stamps off")
