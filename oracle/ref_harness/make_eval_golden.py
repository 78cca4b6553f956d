"""Generates `tests/golden/eval/eval_log.pkl.gz` + `eval_run.npz` by running the UNMODIFIED reference with
`is_evaluating = True` (build container only):

    python oracle/ref_harness/make_eval_golden.py

TEST INFRASTRUCTURE ONLY.  The pickle is what `CattleAviary.evaluation_save()` writes
(`utils/evaluation.py:73-94`, fed by `BaseAviary.update_evaluation_metrics` :1406-1435 and
`evaluation_episode_trigger` :1439-1450); `eval_run.npz` holds what a replay needs: the action of every step, the full
simulator state after every reset, and which steps were followed by a reset.  The run is laid out so that every ordering
quirk of the reference's log shows up in the file:

  episode 0  zero actions (hover) up to the time limit -> trigger; the trigger runs inside `_computeTruncated`, BEFORE
             that step's `update_evaluation_metrics`, so the truncating step's sample opens the next logged episode
  episode 1  uniform random actions until a safety truncation (no evaluation trigger: its samples STAY in the
             evaluator's per-step buffers, which only `append_episode_data` clears)
  episode 2  hover to the time limit again -> second trigger, whose per-step lists therefore hold three episodes' samples
             (the PID memory survives resets, SURVEY fact 8, and a hover with the integrals of episode 1 tips over within
             14 steps -- so the harness clears the controllers through their own `reset()` before this episode)
  episode 3  a few steps that are never logged at episode level
"""
import contextlib
import gzip
import io
import os
import pickle
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_env  # noqa: E402

OUT = os.path.abspath(os.path.join(HERE, "..", "..", "tests", "golden", "eval"))
LEVEL = 7       # the level the reference hard-codes (CattleAviary.py:62): EPISODE_LEN_SEC = 80


def main():
    table = ref_env.synthetic_spawn_table(100, 20, seed=0)
    env = ref_env.make_reference_env(3, 20, ctrl_freq=30, level=LEVEL, spawn_table=table, seed=11)
    env.is_evaluating = True
    rng = np.random.default_rng(5)
    plan = ["hover", "random", "hover", "random"]
    actions, reset_after, reset_states, flags = [], [], [], []
    limit = int(env.EPISODE_LEN_SEC * env.CTRL_FREQ / env.PYB_STEPS_PER_CTRL) + 3
    for ep, kind in enumerate(plan):
        ref_env.quiet(env.reset)
        if kind == "hover":
            for c in env.ctrl:
                c.reset()
        reset_states.append(ref_env.capture_state(env))
        for t in range(limit if kind == "hover" else (6 if ep == len(plan) - 1 else 200)):
            a = np.zeros((12, 4), dtype=np.float32) if kind == "hover" else rng.uniform(-1, 1, (12, 4)).astype(np.float32)
            obs, rew, term, trunc, _ = ref_env.quiet(env.step, a[: env.NUM_DRONES])
            actions.append(a)
            flags.append((bool(term), bool(trunc)))
            done = bool(term) or bool(trunc)
            reset_after.append(done)
            if done:
                break
        print(f"episode {ep} ({kind}): {t + 1} steps, last flags {flags[-1]}")
    path = os.path.join(OUT, "eval_log.pkl")
    os.makedirs(OUT, exist_ok=True)
    with contextlib.redirect_stdout(io.StringIO()):
        env.eval_system.save_evaluation_data(path)      # the reference's own writer (utils/evaluation.py:73-94)
    with open(path, "rb") as f:
        raw = f.read()
    d = pickle.loads(raw)
    with gzip.open(path + ".gz", "wb", compresslevel=9) as f:     # the herd does not move on the DYN path: compresses 30x
        f.write(raw)
    os.remove(path)
    print("logged episodes:", len(d["num_drones"]), "steps per logged episode:", [len(x) for x in d["time_per_step"]],
          "time_taken:", d["time_taken"])
    rec = {"action": np.stack(actions), "reset_after": np.array(reset_after, dtype=np.uint8),
           "terminated": np.array([f[0] for f in flags], dtype=np.uint8), "truncated": np.array([f[1] for f in flags], dtype=np.uint8),
           "spawn_table": table, "level": np.int32(LEVEL)}
    for k in reset_states[0]:
        rec["reset_" + k] = np.stack([s[k] for s in reset_states])
    np.savez_compressed(os.path.join(OUT, "eval_run.npz"), **rec)
    print("wrote", OUT, {f: os.path.getsize(os.path.join(OUT, f)) // 1024 for f in os.listdir(OUT)}, "KiB")


if __name__ == "__main__":
    main()
