"""Herd spawn tables: the reset-time input data of the reference.

`load_yaml` reads a file in the format of the reference's `config/cattle_positions.yaml`
(`simulations[i].cows[j].{x,y}`, read at sb3_envs/BaseAviary.py:88-94).  `generate` follows the algorithm of
the reference's generator script (`utils/cattle_spawn.py:4-46`: herd centre on an 8..13 m annulus, cows at
uniform offsets with a 0.8 m minimum spacing, coordinates rounded to 3 decimals); the shipped table has 16
cows per layout, so BASELINE.json's 20- and 100-cow configurations need a generated one.
"""
import numpy as np


def load_yaml(path):
    import yaml
    with open(path) as f:
        d = yaml.safe_load(f)
    return np.array([[[c["x"], c["y"]] for c in sim["cows"]] for sim in d["simulations"]], dtype=np.float64)


def generate(n_sims, n_cows, seed=0, min_radius=8.0, max_radius=13.0, min_cow_distance=0.8):
    """Returns float64 (n_sims, n_cows, 2).  The offset half-range is widened from the script's 2 m to
    max(2, 0.5*sqrt(n_cows)) m so that the minimum spacing stays feasible for large herds."""
    rng = np.random.RandomState(seed)
    half = max(2.0, 0.5 * np.sqrt(n_cows))
    out = np.zeros((n_sims, n_cows, 2))
    for s in range(n_sims):
        theta = rng.uniform(0, 2 * np.pi)
        r = rng.uniform(min_radius, max_radius)
        centre = np.array([r * np.cos(theta), r * np.sin(theta)])
        placed = np.zeros((0, 2))
        for c in range(n_cows):
            while True:
                off = rng.uniform(-half, half, size=2)
                off *= rng.choice([-1, 1], size=2)
                pos = centre + off
                if placed.shape[0] == 0 or np.all(np.linalg.norm(placed - pos, axis=1) >= min_cow_distance):
                    placed = np.vstack([placed, pos])
                    out[s, c] = np.round(pos, 3)
                    break
    return out
