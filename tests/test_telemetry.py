"""Evaluation telemetry: the reference's pickle schema (utils/evaluation.py:73-94) from the GPU simulator."""
import os
import pickle
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import rl_cattle_herding_b200 as pkg  # noqa: E402
from rl_cattle_herding_b200 import telemetry  # noqa: E402


def test_effectiveness_matches_oracle_and_reference():
    """Winding-number effectiveness against the oracle's restatement and, when the reference tree is mounted (build
    container only), against `utils/evaluation.evaluate_herding_effectiveness` itself."""
    import cattle_oracle as co
    rng = np.random.default_rng(3)
    ref_fn = None
    ref_file = "/root/reference/gym_pybullet_drones/utils/evaluation.py"
    if os.path.exists(ref_file):
        import importlib.util
        spec = importlib.util.spec_from_file_location("_ref_eval", ref_file)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        ref_fn = mod.evaluate_herding_effectiveness
    for trial in range(40):
        n, m = int(rng.integers(3, 9)), int(rng.integers(1, 30))
        drones = rng.uniform(-4, 4, (n, 2))
        cows = rng.uniform(-3, 3, (m, 2))
        o = co.OracleEnv(m, spawn_table=np.zeros((1, m, 2)))
        o.n_drones = n
        o.pos[:n, :2] = drones
        o.cattle_pos[:, :2] = cows
        got = telemetry.herding_effectiveness(cows, drones)
        assert got == pytest.approx(o.effectiveness(), abs=1e-12)
        if ref_fn is not None:
            assert got == pytest.approx(ref_fn(cows, drones), abs=1e-12)


def test_schema_keys():
    assert telemetry.KEYS == ("distances", "num_drones", "time_taken", "effectiveness", "distances_per_step", "time_per_step",
                              "effectiveness_per_step", "drone_poses_per_step", "cattle_poses_per_step", "drone_vel_per_step",
                              "cattle_vel_per_step")


@pytest.mark.gpu
def test_evaluation_pickle_from_single_env_facade(tmp_path):
    """CattleAviary facade with is_evaluating: hover to the time limit of a short level, save, read the pickle back."""
    # float64 build: a perfectly symmetric hover stays a hover (the fp32 build leaves it after ~25 steps)
    env = pkg.CattleAviary(num_drones=12, num_cattle=16, ctrl_freq=30, fixed_num_drones=3, curriculum_level=0, precision="fp64")
    env.is_evaluating = True
    env.reset()
    limit = int(env.EPISODE_LEN_SEC * env.CTRL_FREQ / env.PYB_STEPS_PER_CTRL) + 2     # the limit counts physics steps
    steps = 0
    for _ in range(limit + 5):
        obs, rew, term, trunc, info = env.step(np.zeros((3, 4), dtype=np.float32))
        steps += 1
        if term or trunc:
            break
    assert trunc and not term
    path = str(tmp_path / "evaluation_data.pkl")
    env.evaluation_save(path)
    with open(path, "rb") as f:
        d = pickle.load(f)
    assert tuple(d.keys()) == telemetry.KEYS
    assert len(d["num_drones"]) == 1 and d["num_drones"][0] == 3 and len(d["distances"][0]) == 3
    assert len(d["time_per_step"][0]) == steps
    assert d["time_taken"][0] == pytest.approx(steps * env.PYB_STEPS_PER_CTRL / env.CTRL_FREQ)
    assert np.asarray(d["drone_poses_per_step"][0][0]).shape == (3, 2) and np.asarray(d["cattle_poses_per_step"][0][0]).shape == (16, 2)
    # hover: drones stay put, so every later step adds (almost) nothing to the 1.7 x path length; the first step is
    # measured from the origin like the reference (last_drones_pos is zeroed by reset)
    d0 = np.asarray(d["distances_per_step"][0][0])
    spawn_x = np.array([0.0, 1.75, 3.5])
    assert np.allclose(d0[:, 0], spawn_x + 1.7 * spawn_x, atol=1e-3)
    dl = np.asarray(d["distances_per_step"][0][-1])
    assert np.all(dl >= d0 - 1e-6) and np.all(dl - d0 < 0.5)
    assert all(0.0 <= x <= 100.0 for x in d["effectiveness_per_step"][0])
    env.close()
