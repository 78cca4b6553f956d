"""Stand-in for `pybullet_data` (TEST INFRASTRUCTURE ONLY): only getDataPath() is used
(`sb3_envs/BaseAviary.py:587`)."""


def getDataPath():
    return ""
