"""Per-phase latency of the fused step kernel from in-kernel clock64() stamps (GPU box).  The stamps live in the FULL build;
to time the lean build that ships, point CHS_LIB at a library compiled with -DCHS_PROFILE_LEAN.
    python tools/phase_timing.py [envs] [G] [threads] [drones] [cattle]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import rl_cattle_herding_b200 as pkg  # noqa: E402

E = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
G = int(sys.argv[2]) if len(sys.argv) > 2 else 8
TH = int(sys.argv[3]) if len(sys.argv) > 3 else 128
N = int(sys.argv[4]) if len(sys.argv) > 4 else 3
M = int(sys.argv[5]) if len(sys.argv) > 5 else 20
env = pkg.CattleHerdingBatchedEnv(E, num_drones=N, num_cattle=M, fixed_num_drones=N, spawn_table=pkg.spawn.generate(100, M),
                                  device=0, envs_per_block=G, threads_per_block=TH)
env.reset()
grid = (E + G - 1) // G
clk = torch.zeros((grid, 64), dtype=torch.int64, device="cuda")
acts = [torch.rand((E, N, 4), device="cuda") * 2 - 1 for _ in range(8)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for i in range(40):
    env.step(acts[i % 8])
pkg._lib.check(env.lib.chs_set_phase_clock(env._h, C.c_void_p(clk.data_ptr())))
acc = np.zeros(12)
tot = 0.0
R = 20
for i in range(R):
    if not os.environ.get("NO_FLUSH"):
        flush.zero_()
    env.step(acts[i % 8])
    torch.cuda.synchronize()
    c = clk.cpu().numpy().astype(np.float64)
    d = np.diff(c[:, :13], axis=1)
    acc += d.mean(axis=0)
    tot += 0
names = ["p0 issue+fill", "p0 barrier", "p1 thread0 chain", "p1 barrier", "p2 thread0 work", "p2 barrier", "p3 reduce+cent",
         "p3 reward/flags", "p3 stores+stats", "p3 barrier", "p4 copy-out", "p5 reset"]
print(f"E={E} G={G} threads={TH} grid={grid}: mean cycles per phase (avg over CTAs, {R} steps)")
for n, v in zip(names, acc / R):
    print(f"  {n:18s} {v:9.0f} cyc")
print(f"  sum {acc.sum() / R:.0f} cyc")
c = clk.cpu().numpy().astype(np.float64)
print("  last step, p3 detail: barrier->loop %.0f, loop %.0f, after loop->stamp7 %.0f" % ((c[:, 13] - c[:, 6]).mean(), (c[:, 14] - c[:, 13]).mean(), (c[:, 7] - c[:, 14]).mean()))

def span(a, b):
    return (c[:, b] - c[:, a]).mean()
print("  p1 detail: start->action/PID done %.0f, substeps+euler %.0f, stores %.0f" % (span(2, 16), span(16, 17), span(17, 3)))
print("  p2 detail: neighbours %.0f, cattle obs loop %.0f, herd finish %.0f, winding %.0f" % (span(4, 19), span(19, 20), span(20, 21), span(21, 5)))
print("  p2 neighbours detail: ranking loop %.0f, obs slots %.0f, spacing rewards %.0f, tail %.0f" % (span(4, 24), span(24, 25), span(25, 26), span(26, 19)))
print("  p3 detail: reward->stats done %.0f, env stores %.0f, bulk issue %.0f" % (span(8, 22), span(22, 23), span(23, 9)))
tot = c[:, 12] - c[:, 0]
p5 = c[:, 12] - c[:, 11]
p3s = c[:, 9] - c[:, 8]
busy = p5 > np.median(p5) * 1.5 + 200
print("  last step: CTA total cycles min/median/p90/max = %.0f / %.0f / %.0f / %.0f" % (tot.min(), np.median(tot), np.percentile(tot, 90), tot.max()))
print("  CTAs with a long p5 (reset): %d of %d; p5 %.0f vs %.0f cyc; p3 stores+stats %.0f vs %.0f; total %.0f vs %.0f" % (
    busy.sum(), len(busy), p5[busy].mean() if busy.any() else 0, p5[~busy].mean(), p3s[busy].mean() if busy.any() else 0, p3s[~busy].mean(),
    tot[busy].mean() if busy.any() else 0, tot[~busy].mean()))
g0, g1 = c[:, 30], c[:, 31]
t0 = g0.min()
print("  last step, global timer [us] from the first CTA's start: CTA starts median %.2f / max %.2f; CTA ends min %.2f / median %.2f / max %.2f; CTA duration median %.2f" % (
    np.median(g0 - t0) / 1e3, (g0 - t0).max() / 1e3, (g1 - t0).min() / 1e3, np.median(g1 - t0) / 1e3, (g1 - t0).max() / 1e3, np.median(g1 - g0) / 1e3))

print("  since stamp 4: before s.n load %.0f, after it %.0f, drone_neighbors entry %.0f, after me load %.0f" % tuple((c[:, k] - c[:, 4]).mean() for k in (41, 42, 40, 41 - 41 + 24 + 17)))
print("  ranking loop stamps (cycles since stamp 4): entry %.0f; per iteration: %s" % ((c[:, 31] - c[:, 4]).mean(), " ".join("%.0f" % (c[:, 32 + j] - c[:, 4]).mean() for j in range(N))))
