"""B200-native batched simulator for the RL-Cattle-Herding env step (Physics.DYN path).

The directory name carries a hyphen (task layout); import it as `rl_cattle_herding_b200` through the shim
module of that name at the repository root.
"""
from . import _lib  # noqa: F401
from .enums import ActionType, DroneModel, ImageType, ObservationType, Physics  # noqa: F401
from . import spawn  # noqa: F401


def __getattr__(name):
    # torch-dependent layers are imported lazily so that ABI/host-logic tests stay light
    if name == "CattleHerdingBatchedEnv":
        from .batched_env import CattleHerdingBatchedEnv
        return CattleHerdingBatchedEnv
    if name == "GpuCattleVecEnv":
        from .vec_env import GpuCattleVecEnv
        return GpuCattleVecEnv
    if name in ("BaseAviary", "BaseRLAviary", "CattleAviary"):
        from . import aviary
        return getattr(aviary, name)
    if name in ("MARLCattleAviary", "RLlibMultiAgentWrapper", "GpuMARLBatchedEnv"):
        from . import marl_env
        return getattr(marl_env, name)
    raise AttributeError(name)
