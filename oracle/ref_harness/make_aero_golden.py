"""Generates `tests/golden/aero/aero_forces.npz`: force vectors computed by the UNMODIFIED reference's aerodynamic models
(build container only).

    python oracle/ref_harness/make_aero_golden.py

`_groundEffect`, `_drag` and `_downwash` (sb3_envs/BaseAviary.py:943-1039) are plain Python that ends in
`p.applyExternalForce(...)`; the stand-in pybullet records those calls (standins/pybullet.py: force_log) and serves
`getLinkStates` from the URDF's propeller offsets.  For every sample the drones' kinematic state is written into the
reference env (its cached arrays AND the body store), the three methods are called on drone 0 and the recorded forces are
stored next to the inputs.  Samples cover low altitude (clipped propeller height), |roll| or |pitch| beyond pi/2 (the
ground-effect gate), stacked drones (downwash), drones farther than 10 m and drones below.
`tests/test_oracle_extras.py` holds `oracle/cattle_oracle.py: aero_terms` to these values at 1e-12, and the CUDA tests
compare the device against that restatement.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_env  # noqa: E402

OUT = os.path.abspath(os.path.join(HERE, "..", "..", "tests", "golden", "aero", "aero_forces.npz"))


def main(n_samples=240, seed=7):
    env = ref_env.make_reference_env(4, 16, ctrl_freq=30)
    ref_env.quiet(env.reset)
    import pybullet as p
    rng = np.random.default_rng(seed)
    N = env.NUM_DRONES
    rec = {k: [] for k in ("pos", "quat", "rpy", "vel", "rpm", "drag_rpm", "gnd_forces", "gnd_applied", "drag_force",
                           "downwash_forces")}
    for s in range(n_samples):
        pos = rng.uniform(-2, 2, (N, 3))
        pos[:, 2] = rng.uniform(0.02, 1.5, N)
        kind = s % 6
        if kind == 0:                       # very low: clipped propeller heights
            pos[0, 2] = rng.uniform(0.0, 0.06)
        if kind in (1, 2):                  # stacked: others above drone 0 within a few decimetres
            pos[1:, :2] = pos[0, :2] + rng.normal(0, 0.08, (N - 1, 2))
            pos[1:, 2] = pos[0, 2] + rng.uniform(0.05, 1.2, N - 1)
        if kind == 3:                       # one far away (> 10 m), one below
            pos[1, :2] = pos[0, :2] + np.array([11.0, 3.0])
            pos[1, 2] = pos[0, 2] + 0.4
            pos[2, 2] = max(0.01, pos[0, 2] - 0.3)
        rpy = rng.uniform(-0.4, 0.4, (N, 3))
        if kind == 4:                       # beyond the ground-effect gate
            rpy[0, rng.integers(0, 2)] = rng.choice([-1, 1]) * rng.uniform(np.pi / 2 + 0.05, 2.8)
        if kind == 5:
            rpy[0] = rng.uniform(-1.5, 1.5, 3)
        rpy[:, 2] = rng.uniform(-np.pi, np.pi, N)
        vel = rng.uniform(-3, 3, (N, 3))
        rpm = rng.uniform(9000, 21700, 4)
        drag_rpm = rng.uniform(9000, 21700, 4)
        quat = np.array([p.getQuaternionFromEuler(rpy[i]) for i in range(N)])
        for i in range(N):
            p.resetBasePositionAndOrientation(env.DRONE_IDS[i], pos[i], quat[i], physicsClientId=env.CLIENT)
            p.resetBaseVelocity(env.DRONE_IDS[i], vel[i], [0, 0, 0], physicsClientId=env.CLIENT)
        env._updateAndStoreKinematicInformation()
        st_pos, st_quat, st_rpy, st_vel = env.pos[:N].copy(), env.quat[:N].copy(), env.rpy[:N].copy(), env.vel[:N].copy()

        def record(fn, *a):
            del p.force_log[:]
            fn(*a)
            return list(p.force_log)

        g = record(env._groundEffect, rpm, 0)
        gf = np.zeros(4)
        for c in g:
            assert c["kind"] == "force" and c["flags"] == p.LINK_FRAME and c["force"][0] == 0 and c["force"][1] == 0
            gf[c["link"]] = c["force"][2]
        d = record(env._drag, drag_rpm, 0)
        assert len(d) == 1 and d[0]["link"] == 4 and d[0]["flags"] == p.LINK_FRAME
        w = record(env._downwash, 0)
        wz = np.zeros(N)
        for j, c in enumerate(w):
            assert c["link"] == 4 and c["force"][0] == 0 and c["force"][1] == 0
            wz[j] = c["force"][2]
        rec["pos"].append(st_pos); rec["quat"].append(st_quat); rec["rpy"].append(st_rpy); rec["vel"].append(st_vel)
        rec["rpm"].append(rpm); rec["drag_rpm"].append(drag_rpm)
        rec["gnd_forces"].append(gf); rec["gnd_applied"].append(np.uint8(len(g) == 4))
        rec["drag_force"].append(np.array(d[0]["force"])); rec["downwash_forces"].append(wz)
    out = {k: np.stack(v) for k, v in rec.items()}
    out["consts"] = np.array([env.KF, env.GND_EFF_COEFF, env.PROP_RADIUS, env.GND_EFF_H_CLIP, env.DRAG_COEFF[0], env.DRAG_COEFF[2],
                              env.DW_COEFF_1, env.DW_COEFF_2, env.DW_COEFF_3])
    np.savez_compressed(OUT, **out)
    print(f"{OUT}: {n_samples} samples; ground effect applied in {int(out['gnd_applied'].sum())}, "
          f"downwash contributions {int((out['downwash_forces'] != 0).sum())}, {os.path.getsize(OUT) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
