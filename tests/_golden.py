"""Shared helpers: load `tests/golden/*.npz` (recorded from the reference, see
`oracle/ref_harness/make_golden.py`) and drive the CPU oracle with them."""
import glob
import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
if os.path.join(ROOT, "oracle") not in sys.path:
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import cattle_oracle as co  # noqa: E402

STATE_FIELDS = co.STATE_FIELDS
FLOAT_STATE = ("pos", "quat", "rpy", "vel", "ang_v", "rpy_rates", "last_rpm", "pid_ipos", "pid_last_rpy",
               "pid_irpy", "cattle_pos", "cattle_vel", "prev_cent_dist", "spacing_clock")
INT_STATE = ("n_drones", "step_counter", "step_counter_A", "level", "tally", "spawn_idx")


def golden_names(packing="ctde"):
    """Fixture names for one packing: `m_*` files hold MARL (rllib_envs) rollouts, the rest CTDE (sb3_envs)."""
    names = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
    marl = [n for n in names if n.startswith("m_")]
    return marl if packing == "marl" else [n for n in names if n not in marl]


class Golden:
    def __init__(self, name):
        self.name = name
        z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        self.z = {k: z[k] for k in z.files}
        self.meta = json.loads(str(self.z["meta_json"]))
        self.steps = self.z["action"].shape[0]
        self.table = self.z["spawn_table"]

    def state(self, t, which="before"):
        return {k: self.z[f"{which}_{k}"][t] for k in STATE_FIELDS}

    def make_oracle(self):
        m = self.meta
        return co.OracleEnv(m["n_cattle"], ctrl_freq=m["ctrl_freq"], pyb_freq=m["pyb_freq"],
                            start_level=m["start_level"], spawn_table=self.table, act=m["act"])

    def action(self, t):
        a = self.z["action"][t]
        return a


def max_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    same_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    d = np.where(both_nan | same_inf, 0.0, np.abs(a - b))
    return float(np.max(d)) if d.size else 0.0


def near_tie_rows(pos, n, tol=2e-6):
    """Rows (drones) whose neighbour ranking is decided by a distance gap below fp32 resolution.

    `_computeObs` ranks the other drones by xy distance (BaseRLAviary.py:310-311).  The reference's line
    spawn puts drone 1 exactly between drones 0 and 2, and ring formations are symmetric, so the top-2 order
    can hinge on differences of ~1e-10 m that a float32 state cannot represent.  For such rows the fp32 build
    may legitimately swap equidistant neighbours; the tests skip the neighbour columns of those rows only.
    """
    pos = np.asarray(pos, dtype=np.float64)[:n, :2]
    out = np.zeros((pos.shape[0],), dtype=bool)
    for i in range(n):
        d = np.sort(np.delete(np.sqrt(np.sum((pos - pos[i]) ** 2, axis=1)), i))
        gaps = np.diff(d[:3])
        scale = np.maximum(1.0, np.abs(pos).max())
        out[i] = bool(np.any(gaps < tol * scale))
    return out
