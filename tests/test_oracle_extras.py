"""CPU: sanity of the non-reference extensions in the oracle (aerodynamic extras, kinematic herd)."""
import numpy as np

from _golden import Golden

import cattle_oracle as co  # noqa: E402  (tests/_golden.py put oracle/ on sys.path)


def test_aero_physics_sanity_of_the_restatement():
    """Signs of the three models in the restatement that defines them (there is no Bullet oracle for DYN + extras)."""
    table = np.zeros((2, 16, 2)); table[:, :, 0] = np.arange(16) + 5.0
    hover = np.full((2, 4), co.HOVER_RPM)

    def run(flags, setup):
        o = co.OracleEnv(16, ctrl_freq=30, spawn_table=table, aero_flags=flags)
        o.reset(2, np.full((16, 2), 0.5))
        setup(o)
        pos0 = o.pos.copy()
        for sub in range(8):
            for k in range(2):
                o._dynamics(k, hover[k], hover[k], pos0)
        return o
    def low(o): o.pos[:2, 2] = 0.05
    assert run(2, low).vel[0, 2] > run(0, low).vel[0, 2] + 1e-3          # ground effect lifts a low drone
    def fast(o): o.vel[0] = [2.0, 0, 0]
    assert 0 < run(1, fast).vel[0, 0] < run(0, fast).vel[0, 0]            # drag slows it down
    def stacked(o): o.pos[0] = [0, 0, 0.4]; o.pos[1] = [0, 0, 0.9]
    assert run(4, stacked).vel[0, 2] < run(0, stacked).vel[0, 2] - 1e-3   # downwash pushes the lower drone down
    assert abs(run(4, stacked).vel[1, 2] - run(0, stacked).vel[1, 2]) < 1e-12   # and leaves the upper one alone



def test_extras_off_is_the_reference_path():
    """aero_flags = 0 / cattle_integrate = False leave the pinned DYN path untouched; with cattle_integrate the
    herd advances by vel * CTRL_TIMESTEP per step and nothing else changes on that step."""
    g = Golden("a_3d20c_uniform")
    a, b = g.make_oracle(), g.make_oracle()
    b.cattle_integrate = True
    a.set_state(g.state(0, "before")); b.set_state(g.state(0, "before"))
    a.step(g.action(0)); b.step(g.action(0))
    assert np.array_equal(a.pos, b.pos) and np.array_equal(a.cattle_vel, b.cattle_vel)
    assert np.allclose(b.cattle_pos[:, :2], a.cattle_pos[:, :2] + a.cattle_vel[:, :2] / g.meta["ctrl_freq"], atol=1e-15)
