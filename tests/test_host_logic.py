"""CPU: host-side logic -- spawn tables, the Philox mirror, enums, bench arithmetic."""
import os
import sys

import numpy as np

from _golden import Golden

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def _pkg():
    import rl_cattle_herding_b200 as pkg
    return pkg


def test_spawn_generator_reproduces_the_fixture_tables():
    """`spawn.generate` follows utils/cattle_spawn.py; the fixtures carry the tables the reference was run with."""
    pkg = _pkg()
    g = Golden("a_3d20c_uniform")
    assert np.array_equal(pkg.spawn.generate(100, 20, seed=0), g.table)
    c = Golden("c_8d100c_uniform")
    assert np.array_equal(pkg.spawn.generate(8, 100, seed=0), c.table)
    t = pkg.spawn.generate(3, 20, seed=5)
    for s in range(3):   # 0.8 m minimum spacing of the generator (rounding to 3 decimals may shave 1e-3)
        d = np.linalg.norm(t[s][:, None] - t[s][None], axis=-1) + np.eye(20) * 10
        assert d.min() >= 0.8 - 2e-3


def test_spawn_yaml_loader(tmp_path):
    import yaml
    pkg = _pkg()
    t = pkg.spawn.generate(2, 16, seed=1)
    doc = {"simulations": [{"id": s + 1, "cows": [{"id": c + 1, "x": float(t[s, c, 0]), "y": float(t[s, c, 1])}
                                                   for c in range(16)]} for s in range(2)]}
    p = tmp_path / "cattle_positions.yaml"
    p.write_text(yaml.dump(doc, sort_keys=False))
    assert np.array_equal(pkg.spawn.load_yaml(str(p)), t)


def test_philox_known_answers():
    from rl_cattle_herding_b200 import philox as ph
    # Random123 known-answer vectors for philox4x32-10
    assert [int(x) for x in ph.philox4x32_10(0, 0, 0, 0, 0)] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert [int(x) for x in ph.philox4x32_10(0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffffffffffff)] == \
        [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    u = ph.cow_uniforms(7, 3, 1, 20)
    assert u.shape == (20, 2) and (u >= 0).all() and (u < 1).all()
    assert not np.allclose(u, ph.cow_uniforms(7, 4, 1, 20))       # keyed by env id
    assert not np.allclose(u, ph.cow_uniforms(7, 3, 2, 20))       # ... and by episode
    ns = [ph.num_drones(1, e, 1, 4, 12) for e in range(400)]
    assert min(ns) == 4 and max(ns) == 12


def test_enums_match_reference_values():
    pkg = _pkg()
    assert pkg.Physics.DYN.value == "dyn" and pkg.ActionType.VEL.value == "vel"
    assert pkg.ObservationType.COKIN.value == "cokin" and pkg.DroneModel.CF2X.value == "cf2x"


def test_algorithmic_bytes_formula():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.algorithmic_bytes_per_env_step(3, 20, 12, 66) == 4230
    assert bench.algorithmic_bytes_per_env_step(8, 100, 12, 66) == 6870     # config C
