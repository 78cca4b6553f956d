"""Prints the teacher-forced error of the CUDA step against every golden fixture, per field, for both builds.
Run on the GPU box:  python tools/parity_report.py > gpurun_out/parity_report.txt"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
from _golden import FLOAT_STATE, INT_STATE, Golden, golden_names, max_err  # noqa: E402
import rl_cattle_herding_b200 as pkg  # noqa: E402

for precision in ("fp64", "fp32"):
    worst = {}
    for name in golden_names():
        g = Golden(name)
        m = g.meta
        env = pkg.CattleHerdingBatchedEnv(g.steps, num_drones=12, num_cattle=m["n_cattle"], ctrl_freq=m["ctrl_freq"],
                                          pyb_freq=m["pyb_freq"], act=m["act"], start_level=m["start_level"],
                                          auto_reset=False, precision=precision, spawn_table=g.table, device=0)
        env.set_state({k: g.z[f"before_{k}"] for k in FLOAT_STATE + INT_STATE})
        obs, rew, term, trunc = env.step(torch.from_numpy(np.ascontiguousarray(g.z["action"])).cuda())
        torch.cuda.synchronize()
        got = env.get_state()
        line = []
        for k in FLOAT_STATE:
            ref = g.z[f"after_{k}"]
            scale = max(1.0, float(np.max(np.abs(np.where(np.isfinite(ref), ref, 0.0)))))
            e = max_err(got[k], ref)
            worst[k] = max(worst.get(k, 0.0), e / scale)
            line.append(f"{k}={e:.1e}")
        eo = max_err(obs.cpu().numpy(), g.z["obs"])
        rr = g.z["reward"]; fin = np.isfinite(rr)
        er = max_err(rew.cpu().numpy().astype(np.float64)[fin], rr[fin])
        ints = all(np.array_equal(got[k].astype(np.int64), g.z[f"after_{k}"].astype(np.int64)) for k in INT_STATE)
        flags = np.array_equal(term.cpu().numpy(), g.z["terminated"]) and np.array_equal(trunc.cpu().numpy(), g.z["truncated"])
        worst["obs"] = max(worst.get("obs", 0.0), eo); worst["reward"] = max(worst.get("reward", 0.0), er)
        print(f"[{precision}] {name}: ints={'ok' if ints else 'MISMATCH'} flags={'ok' if flags else 'MISMATCH'} "
              f"obs={eo:.1e} reward={er:.1e} " + " ".join(line))
        env.close()
    print(f"[{precision}] WORST scaled: " + " ".join(f"{k}={v:.1e}" for k, v in worst.items()))
