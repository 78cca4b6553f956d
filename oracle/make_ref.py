#!/usr/bin/env python
"""Stages the UNMODIFIED reference package where the GPU box can see it.  TEST / BASELINE INFRASTRUCTURE ONLY.

    python oracle/make_ref.py            # /root/reference/gym_pybullet_drones -> oracle/_ref/gym_pybullet_drones

The reference is pure Python and `/root/reference` does not exist on the GPU box, so the CPU
arm of `bench.py` (`--impl reference`, `cpu_baseline`) could only time a port there.  This recipe copies the files of
the reference's env stack -- byte for byte, nothing edited, sha256 of every file in `MANIFEST.json` -- into the
git-ignored `oracle/_ref/` (it travels with the gpurun snapshot exactly like the built `libchs.so`, and stays out of
the history).  `oracle/ref_harness/ref_env.py` imports it from there when `/root/reference` is absent, together with
the committed stand-ins for the packages that are installed on neither box (`oracle/ref_harness/standins`: pybullet as
a body-state store + quaternion helpers, pybullet_data, gymnasium).

Copied: the package's `__init__.py`, `control/`, `config/`, `utils/`, `sb3_envs/`, `rllib_envs/` and the URDF files of
`assets/` (the constants `BaseAviary._parseURDFParameters` reads).  Not copied: `simulator/` (drivers and 230 MB of
saved models), meshes, images.  Nothing under `oracle/_ref` is imported by the product path (tests/test_abi.py).
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/gym_pybullet_drones"
DST_ROOT = os.path.join(HERE, "_ref")
DST = os.path.join(DST_ROOT, "gym_pybullet_drones")
DIRS = ("control", "config", "utils", "sb3_envs", "rllib_envs")
KEEP_EXT = (".py", ".yaml", ".yml", ".urdf")


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def stage(force=False):
    """Returns the staged root (`oracle/_ref`) or None when the reference tree is not on this machine."""
    if not os.path.isdir(SRC):
        return DST_ROOT if os.path.isdir(DST) else None
    manifest_path = os.path.join(DST_ROOT, "MANIFEST.json")
    if os.path.exists(manifest_path) and not force:
        return DST_ROOT
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(DST, exist_ok=True)
    files = {}

    def copy(rel):
        s, d = os.path.join(SRC, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(s, d)
        files[rel] = _sha(d)

    copy("__init__.py")
    for sub in DIRS:
        for root, _, names in os.walk(os.path.join(SRC, sub)):
            for n in sorted(names):
                if n.endswith(KEEP_EXT):
                    copy(os.path.relpath(os.path.join(root, n), SRC))
    for n in sorted(os.listdir(os.path.join(SRC, "assets"))):
        if n.endswith(".urdf"):
            copy(os.path.join("assets", n))
    with open(manifest_path, "w") as f:
        json.dump({"source": SRC, "note": "unmodified copies; sha256 per file", "files": files}, f, indent=1, sort_keys=True)
    return DST_ROOT


if __name__ == "__main__":
    root = stage(force="--force" in sys.argv)
    print("staged:", root)
