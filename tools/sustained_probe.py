"""Per-step time of a long run at a large batch, with SM clock / power samples: burst vs sustained behaviour."""
import os, subprocess, sys, threading, time
import numpy as np, torch
sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import rl_cattle_herding_b200 as pkg
E = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
D, M = 3, 20
env = pkg.CattleHerdingBatchedEnv(E, num_drones=D, num_cattle=M, ctrl_freq=30, fixed_num_drones=D, spawn_table=pkg.spawn.generate(100, M, seed=0), device=0, seed=1, step_mode="split")
env.set_launch(32, 256)
env.reset()
acts = [torch.rand((E, D, 4), device="cuda") * 2 - 1 for _ in range(4)]
samples, stop = [], False
def smi():
    while not stop:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown", "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True).stdout.strip()
        samples.append((time.perf_counter(), out))
        time.sleep(0.05)
th = threading.Thread(target=smi); th.start()
time.sleep(0.3)
N = 600
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(N)]
t0 = time.perf_counter()
for i, (a, b) in enumerate(ev):
    a.record(); env.step(acts[i % 4]); b.record()
torch.cuda.synchronize()
t1 = time.perf_counter()
stop = True; th.join()
ms = np.array([a.elapsed_time(b) for a, b in ev])
for lo in range(0, N, 50):
    print(f"steps {lo:3d}-{lo + 49:3d}: mean {ms[lo:lo + 50].mean() * 1e3:7.1f} us  (even {ms[lo:lo + 50:2].mean() * 1e3:6.1f} odd {ms[lo + 1:lo + 50:2].mean() * 1e3:6.1f})")
print("wall %.1f ms for %d steps" % ((t1 - t0) * 1e3, N))
for t, s in samples:
    if t0 - 0.3 <= t <= t1 + 0.1: print("  smi @%+.2fs: %s" % (t - t0, s))
