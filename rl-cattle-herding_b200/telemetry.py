"""Evaluation telemetry in the reference's wire format.

`utils/evaluation.py:5-94` keeps per-step and per-episode lists on the host and pickles them as one dict
(`save_evaluation_data`); `utils/pickel_me_this.py` plots from that file.  `EvaluationLog` produces the same dict
(same keys, same nesting: episode -> step -> per-entity arrays) for one watched environment.  The numbers that the
reference accumulates step by step come from the simulator's own device-side accumulators -- the per-drone running path
length (`ChsStateView.path_len`, kept by the step kernel) and the herded-cow count of the step (`chs_set_debug_buffers`)
-- the poses the schema stores per step come from one `chs_get_state` of that env.

Pinned to a pickle written by the unmodified reference (`tests/golden/eval/`, `oracle/ref_harness/make_eval_golden.py`),
including the reference's ordering and aliasing quirks:
  * `update_evaluation_metrics` (sb3_envs/BaseAviary.py:1406-1435) runs BEFORE the step counter advances (:462-464), so
    `time_per_step` / `time_taken` use the pre-increment counter;
  * the running "distance" of a drone is a 2-vector that starts at ZERO (`_housekeeping` reads the position cache before
    the first kinematics refresh, :683-688 vs :322) and grows by 1.7 x the xy step length in both components (:1424);
    the first step is measured from the origin because `last_drones_pos` is zeroed by reset (:317);
  * `evaluation_episode_trigger` (:1439-1450) fires inside `_computeTruncated` on the time limit only
    (CattleAviary.py:545-550), i.e. before that step's per-step append -- the truncating step's sample opens the NEXT
    logged episode -- and twice per step, because `_computeReward` calls `_computeTruncated` as well (:315), so every
    time-limit episode is followed by an empty one;
  * the per-step buffers are only cleared by that trigger: samples of episodes that end any other way stay in them;
  * the log stores the list object `episode_drone_distances` itself, which later steps keep updating in place: all entries
    of one episode alias one list and show its final values (pickle keeps the aliasing);
  * `cattle_vel_per_step` is the cached velocity from before the step's herd update (`_getCowStateVector` reads the
    cache filled at :425, `_flockingStep` writes the store afterwards, :1398-1400).
"""
import pickle

import numpy as np

KEYS = ("distances", "num_drones", "time_taken", "effectiveness", "distances_per_step", "time_per_step",
        "effectiveness_per_step", "drone_poses_per_step", "cattle_poses_per_step", "drone_vel_per_step", "cattle_vel_per_step")


def herding_effectiveness(cattle_xy, drones_xy):
    """`evaluate_herding_effectiveness` (utils/evaluation.py:100-138,271-273): percentage of cows with a non-zero
    winding number w.r.t. the polygon of the drones in index order (host-side copy for the log; the step kernel
    computes the same count on the device for the reward)."""
    cattle_xy, drones_xy = np.asarray(cattle_xy, dtype=float), np.asarray(drones_xy, dtype=float)
    n = len(drones_xy)
    if n == 0 or len(cattle_xy) == 0:
        return 0.0
    inside = 0
    for px, py in cattle_xy:
        wn = 0
        for i in range(n):
            ax, ay = drones_xy[i]
            bx, by = drones_xy[(i + 1) % n]
            il = (bx - ax) * (py - ay) - (px - ax) * (by - ay)
            if ay <= py:
                if by > py and il > 0:
                    wn += 1
            elif by <= py and il < 0:
                wn -= 1
        inside += wn != 0
    return inside / len(cattle_xy) * 100.0


class EvaluationLog:
    """One watched environment of a `CattleHerdingBatchedEnv` (`env_index`), logged like the reference's `evaluator`."""

    def __init__(self, sim, env_index=0):
        self.sim, self.e = sim, int(env_index)
        self.data = {k: [] for k in KEYS}
        self._cur = {k: [] for k in KEYS if k.endswith("_per_step")}      # evaluator.curr_*: cleared by the trigger only
        self._dist, self._cvel, self._n = [], None, 0
        if sim._dbg_herd is None:
            sim.enable_debug_outputs()

    def state(self):
        return self.sim.get_state(self.e, 1)

    def _effectiveness(self):
        """Herded-cow count of the last step, as the step kernel's winding-number pass left it on the device."""
        return int(self.sim._dbg_herd[self.e].item()) / self.sim.num_cattle * 100.0     # utils/evaluation.py:136

    def on_reset(self, s=None):
        """After `reset()` (`_housekeeping` :683-688 and `reset` :317): a NEW distances list, starting at zero."""
        s = self.state() if s is None else s
        self._n = int(s["n_drones"][0])
        self._dist = [np.zeros(2) for _ in range(self._n)]
        self._cvel = s["cattle_vel"][0, :, :2].copy()

    def on_step(self, s=None):
        """After every `step()`, with the state it left (`update_evaluation_metrics`)."""
        s = self.state() if s is None else s
        n = self._n
        if len(self._dist) != n or self._cvel is None:
            self.on_reset(s)
            n = self._n
        path = np.maximum(s["path_len"][0, :n], 0.0)          # device accumulator; < 0 marks "no step yet"
        for i in range(n):
            self._dist[i][:] = path[i]                          # in place: earlier log entries alias this list
        dpos, dvel = s["pos"][0, :n, :2].copy(), s["vel"][0, :n, :2].copy()
        cpos = s["cattle_pos"][0, :, :2].copy()
        c = self._cur
        c["distances_per_step"].append(self._dist)
        c["time_per_step"].append((int(s["step_counter"][0]) - self.sim.substeps) / self.sim.ctrl_freq)
        c["effectiveness_per_step"].append(self._effectiveness())
        c["drone_poses_per_step"].append(dpos)
        c["cattle_poses_per_step"].append(cpos)
        c["drone_vel_per_step"].append(dvel)
        c["cattle_vel_per_step"].append(self._cvel)
        self._cvel = s["cattle_vel"][0, :, :2].copy()

    def on_episode_end(self, s=None):
        """`evaluation_episode_trigger` (:1439-1450); call it BEFORE `on_step` of the truncating step, twice."""
        s = self.state() if s is None else s
        d = self.data
        d["distances"].append(self._dist)
        d["num_drones"].append(self._n)
        d["time_taken"].append((int(s["step_counter"][0]) - self.sim.substeps) / self.sim.ctrl_freq)
        d["effectiveness"].append(self._effectiveness())
        for k, v in self._cur.items():
            d[k].append(v)
        self._cur = {k: [] for k in self._cur}

    def save(self, save_path="evaluation_data.pkl"):
        """`evaluator.save_evaluation_data` (utils/evaluation.py:73-94): same dict, same default file name."""
        with open(save_path, "wb") as f:
            pickle.dump({k: self.data[k] for k in KEYS}, f)
        return save_path
