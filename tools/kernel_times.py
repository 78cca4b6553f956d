"""Per-kernel GPU times of consecutive chs_step calls in a normal (unprofiled-by-ncu) run, through CUPTI via
torch.profiler: what each of the step's kernels costs back to back, warm.  usage: python tools/kernel_times.py E [mode]"""
import os
import sys
import torch
from torch.profiler import profile, ProfilerActivity

sys.path.insert(0, os.path.abspath(os.path.join(os.path.dirname(__file__), "..")))
import rl_cattle_herding_b200 as pkg  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
mode = sys.argv[2] if len(sys.argv) > 2 else "auto"
env = pkg.CattleHerdingBatchedEnv(E, num_drones=3, num_cattle=20, ctrl_freq=30, fixed_num_drones=3,
                                  spawn_table=pkg.spawn.generate(100, 20, seed=0), device=0, seed=1, step_mode=mode)
if E >= 65536:
    env.set_launch(32, 256)
env.reset()
acts = [torch.rand((E, 3, 4), device="cuda") * 2 - 1 for _ in range(4)]
for i in range(60):
    env.step(acts[i % 4])
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(40):
        env.step(acts[i % 4])
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
agg = {}
for e in ev:
    a = agg.setdefault(e.name[:60], [0, 0.0])
    a[0] += 1; a[1] += e.time_range.elapsed_us()
span = ev[-1].time_range.end - ev[0].time_range.start
busy = sum(v[1] for v in agg.values())
print(f"E={E} mode={mode}: 40 steps, span {span / 40:.1f} us/step, kernels busy {busy / 40:.1f} us/step, idle {(span - busy) / 40:.1f} us/step")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {t / n:8.1f} us x {n:3d}  {k}")
