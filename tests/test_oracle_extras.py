"""CPU: the aerodynamic extras of the oracle against force vectors computed by the reference's own formulas, and
sanity of the non-reference kinematic herd."""
import os

import numpy as np

from _golden import Golden

import cattle_oracle as co  # noqa: E402  (tests/_golden.py put oracle/ on sys.path)


def test_aero_terms_match_forces_recorded_from_the_reference():
    """tests/golden/aero/aero_forces.npz holds what the UNMODIFIED `_groundEffect`, `_drag`, `_downwash`
    (sb3_envs/BaseAviary.py:943-1039) handed to `p.applyExternalForce` for 240 random states (recorded through the
    stand-in by oracle/ref_harness/make_aero_golden.py).  `aero_terms`, the piece `_dynamics` adds to the DYN step,
    has to reproduce every vector at 1e-12 (relative to the force scale) -- the CUDA tests then compare the device with
    this restatement."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "aero", "aero_forces.npz"))
    n_applied = n_gated = n_dw = 0
    for s in range(z["pos"].shape[0]):
        pos, quat, vel = z["pos"][s], z["quat"][s], z["vel"][s]
        R = co.quat_to_matrix(tuple(quat[0]))
        gnd, drag_w, dwz = co.aero_terms(7, pos[0], R, vel[0], z["rpm"][s], z["drag_rpm"][s], pos, 0, pos.shape[0])
        ref_g = z["gnd_forces"][s]
        assert np.allclose(gnd, ref_g, rtol=1e-12, atol=1e-18), (s, gnd, ref_g)
        assert bool(z["gnd_applied"][s]) == (R[2][2] > 0)          # the |roll|, |pitch| < pi/2 gate (:970)
        n_applied += int(z["gnd_applied"][s]); n_gated += int(not z["gnd_applied"][s])
        # the reference hands R^T d to LINK_FRAME of the centre-of-mass link: rotate ours into the body frame
        Rm = np.array(R)
        assert np.allclose(Rm.T @ np.array(drag_w), z["drag_force"][s], rtol=1e-12, atol=1e-18), s
        ref_dw = float(np.sum(z["downwash_forces"][s]))
        assert abs(dwz - ref_dw) <= 1e-12 * max(1.0, abs(ref_dw)), (s, dwz, ref_dw)
        n_dw += int(ref_dw != 0)
    assert n_applied > 100 and n_gated > 20 and n_dw > 50
    c = z["consts"]
    assert np.allclose(c, [co.KF, co.GND_EFF_COEFF, co.PROP_RADIUS, co.GND_EFF_H_CLIP, co.DRAG_XY, co.DRAG_Z, co.DW1, co.DW2, co.DW3],
                       rtol=1e-15)


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
