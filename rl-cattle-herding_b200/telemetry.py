"""Evaluation telemetry in the reference's wire format.

`utils/evaluation.py:5-94` keeps per-step and per-episode lists on the host and pickles them as one dict
(`save_evaluation_data`); `utils/pickel_me_this.py` plots from that file.  `EvaluationLog` produces the same dict
(same keys, same nesting: episode -> step -> per-entity arrays) from the simulator's state, for the few envs an
evaluation run watches.  The per-step arithmetic follows `BaseAviary.update_evaluation_metrics`
(sb3_envs/BaseAviary.py:1406-1435) including its quirks: the running "distance" of a drone is a 2-vector that starts
at the drone's spawn position (:683-688) and grows by 1.7 x the xy step length in both components (:1424), and the
first step is measured from the origin because `last_drones_pos` is zeroed by reset (:317).
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
        self._new_episode()

    def _new_episode(self):
        self._cur = {k: [] for k in KEYS if k.endswith("_per_step")}
        self._dist, self._last = None, None

    def _state(self):
        s = self.sim.get_state(self.e, 1)
        n = int(s["n_drones"][0])
        return s, n

    def on_reset(self):
        """After `reset()`: distances start at the spawn positions, `last_drones_pos` at the origin."""
        s, n = self._state()
        self._new_episode()
        self._dist = [s["pos"][0, i, :2].astype(np.float32).copy() for i in range(n)]
        self._last = [np.zeros(2) for _ in range(n)]

    def on_step(self):
        """After every `step()` (`update_evaluation_metrics`)."""
        s, n = self._state()
        if self._dist is None or len(self._dist) != n:
            self.on_reset()
        dpos, dvel = s["pos"][0, :n, :2].copy(), s["vel"][0, :n, :2].copy()
        cpos, cvel = s["cattle_pos"][0, :, :2].copy(), s["cattle_vel"][0, :, :2].copy()
        for i in range(n):
            self._dist[i] = self._dist[i] + np.float32(np.linalg.norm(self._last[i] - dpos[i]) * 1.7)
            self._last[i] = dpos[i].copy()
        t = int(s["step_counter"][0]) / self.sim.ctrl_freq
        c = self._cur
        c["distances_per_step"].append([d.copy() for d in self._dist])
        c["time_per_step"].append(t)
        c["effectiveness_per_step"].append(herding_effectiveness(cpos, dpos))
        c["drone_poses_per_step"].append(dpos)
        c["cattle_poses_per_step"].append(cpos)
        c["drone_vel_per_step"].append(dvel)
        c["cattle_vel_per_step"].append(cvel)

    def on_episode_end(self):
        """`evaluation_episode_trigger` (:1439-1450): the reference fires it when the time limit truncates an episode."""
        s, n = self._state()
        d = self.data
        d["distances"].append([x.copy() for x in (self._dist or [])])
        d["num_drones"].append(n)
        d["time_taken"].append(int(s["step_counter"][0]) / self.sim.ctrl_freq)
        d["effectiveness"].append(herding_effectiveness(s["cattle_pos"][0, :, :2], s["pos"][0, :n, :2]))
        for k, v in self._cur.items():
            d[k].append(v)
        self._new_episode()

    def save(self, save_path="evaluation_data.pkl"):
        """`evaluator.save_evaluation_data` (utils/evaluation.py:73-94): same dict, same default file name."""
        with open(save_path, "wb") as f:
            pickle.dump({k: self.data[k] for k in KEYS}, f)
        return save_path
