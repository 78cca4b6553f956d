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


def _same(a, b, tol, where=""):
    """Recursive comparison of the pickle's nesting: lists (episode -> step -> entity) of arrays / numbers."""
    if isinstance(b, (list, tuple)):
        assert isinstance(a, (list, tuple)) and len(a) == len(b), f"{where}: length {len(a)} != {len(b)}"
        for i, (x, y) in enumerate(zip(a, b)):
            _same(x, y, tol, f"{where}[{i}]")
    else:
        x, y = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
        assert x.shape == y.shape, f"{where}: shape {x.shape} != {y.shape}"
        assert np.allclose(x, y, rtol=0.0, atol=tol), f"{where}: max |diff| {np.max(np.abs(x - y)):.3e}"


@pytest.mark.gpu
def test_evaluation_pickle_matches_reference_recording(tmp_path, precision="fp64", tol=1e-9):
    """The pickle `CattleAviary.evaluation_save()` writes, against the one the UNMODIFIED reference wrote for the same
    run (tests/golden/eval, oracle/ref_harness/make_eval_golden.py): two time-limit episodes with a safety-truncated one
    in between -- same keys, same nesting and lengths (incl. the empty episode after every trigger and the truncating
    step's sample opening the next episode), same values, same aliasing of the distances lists.
    Float64 build: the run is 300 FREE-RUNNING hover steps, and the reference's roll loop is a positive feedback that
    leaves a hover within ~25 steps once fp32 rounding breaks its symmetry -- the fp32 build's
    accumulators are held to the recordings step by step in tests/test_cuda_parity.py::test_teacher_forced instead."""
    import gzip
    gold = os.path.join(ROOT, "tests", "golden", "eval")
    run = np.load(os.path.join(gold, "eval_run.npz"))
    with gzip.open(os.path.join(gold, "eval_log.pkl.gz"), "rb") as f:
        ref = pickle.load(f)
    env = pkg.CattleAviary(num_drones=12, num_cattle=20, ctrl_freq=30, fixed_num_drones=3, curriculum_level=int(run["level"]),
                           precision=precision, spawn_table=run["spawn_table"])
    env.is_evaluating = True
    fields = [k[len("reset_"):] for k in run.files if k.startswith("reset_")]

    def start_episode(ep):
        env.reset()
        st = {k: np.asarray(run["reset_" + k][ep])[None] for k in fields}
        st["path_len"] = -np.ones((1, 12))          # "no step yet", what a reset leaves
        env.set_state(st)

    ep = 0
    start_episode(ep)
    for t in range(run["action"].shape[0]):
        obs, rew, term, trunc, info = env.step(run["action"][t])
        assert (term, trunc) == (bool(run["terminated"][t]), bool(run["truncated"][t])), f"flags differ at step {t}"
        if run["reset_after"][t]:
            ep += 1
            start_episode(ep)
    path = str(tmp_path / "evaluation_data.pkl")
    env.evaluation_save(path)
    with open(path, "rb") as f:
        d = pickle.load(f)
    assert tuple(d.keys()) == tuple(ref.keys()) == telemetry.KEYS
    assert d["num_drones"] == ref["num_drones"]
    assert [len(x) for x in d["time_per_step"]] == [len(x) for x in ref["time_per_step"]] == [301, 0, 320, 0]
    for k in telemetry.KEYS:
        _same(d[k], ref[k], tol if k != "time_taken" and k != "time_per_step" else 1e-12, k)
    # aliasing, as unpickled from the reference's file: every entry of one env episode is ONE list object
    for dd in (ref, d):
        assert dd["distances_per_step"][0][0] is dd["distances"][0] and dd["distances"][1] is dd["distances"][0]
        assert dd["distances_per_step"][2][0] is dd["distances"][0]           # the truncating step of episode 0
        assert dd["distances_per_step"][2][1] is dd["distances_per_step"][2][18]
        assert dd["distances_per_step"][2][19] is dd["distances"][2]
    # the same accumulators, summed over episodes on the device (chs_stats slots 11-13)
    st = env.sim.stats()
    assert st[0] == 3 and st[8] == 2
    per_episode = [np.sum([x[0] for x in ref["distances"][0]]), np.sum([x[0] for x in ref["distances_per_step"][2][1]]),
                   np.sum([x[0] for x in ref["distances"][2]])]
    assert st[11] == pytest.approx(sum(per_episode), abs=10 * tol)
    assert st[12] == pytest.approx(0.0) and st[13] == pytest.approx(sum(sum(x) for x in ref["effectiveness_per_step"]), abs=1e-9)
    env.close()
