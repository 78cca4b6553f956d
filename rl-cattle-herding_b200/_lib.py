"""ctypes binding of the C ABI declared in `include/chs.h` (libchs.so, built by `__graft_entry__.build()`).

There is no CPU fallback: if the shared library is missing or a call fails, this module raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CHS_LIB", os.path.join(_HERE, "libchs.so"))   # CHS_LIB: build-variant experiments

ABI_VERSION = 4
MAX_DRONES = 12
NUM_STATS = 16
STATIC_SCRATCH_BYTES = 8 * 1024 * 1024

STAT_NAMES = ("episodes", "sum_return", "sum_length", "terminated", "trunc_altitude", "trunc_collision",
              "trunc_formation", "trunc_boundary", "trunc_time_limit", "non_finite", "env_steps", "sum_path_length",
              "sum_final_effectiveness", "sum_step_effectiveness", "curriculum_successes", "level_ups")

# every symbol include/chs.h declares (checked by tests/test_abi.py)
SYMBOLS = ("chs_abi_version", "chs_last_error", "chs_create", "chs_destroy", "chs_get_dims", "chs_set_spawn_table",
           "chs_reset", "chs_step", "chs_step_host", "chs_reset_host", "chs_get_state", "chs_set_state",
           "chs_set_debug_buffers", "chs_set_phase_clock", "chs_set_step_mode", "chs_set_register_variant", "chs_stats", "chs_set_launch",
           "chs_step_static_scratch", "chs_get_num_drones")


class ChsParams(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("num_cattle", C.c_int32), ("max_drones", C.c_int32),
                ("fixed_num_drones", C.c_int32), ("pyb_freq", C.c_int32), ("ctrl_freq", C.c_int32),
                ("act_type", C.c_int32), ("start_level", C.c_int32), ("obs_rows", C.c_int32),
                ("packing", C.c_int32), ("precision", C.c_int32), ("auto_reset", C.c_int32),
                ("aero_flags", C.c_int32), ("cattle_integrate", C.c_int32), ("env_id_offset", C.c_int32),
                ("envs_per_block", C.c_int32), ("threads_per_block", C.c_int32), ("step_mode", C.c_int32),
                ("persistent_padding", C.c_int32), ("marl_wrapper_quirks", C.c_int32), ("seed", C.c_uint64)]


class ChsDims(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("num_envs", "max_drones", "num_cattle", "action_width", "obs_rows",
                                         "obs_len", "substeps", "action_buffer_size", "reward_width",
                                         "envs_per_block", "threads_per_block", "smem_bytes", "flag_width")]


_F64 = ("pos", "quat", "rpy", "vel", "ang_v", "rpy_rates", "last_rpm", "pid_ipos", "pid_last_rpy", "pid_irpy",
        "cattle_pos", "cattle_vel", "prev_cent_dist", "spacing_clock", "ep_return")
_I64 = ("step_counter",)
_I32 = ("step_counter_A", "n_drones", "level", "tally", "spawn_idx", "episode")
_I32_TAIL = ("active_agents",)  # appended to ChsStateView in ABI 4
_F64_TAIL = ("path_len",)


class ChsStateView(C.Structure):
    _fields_ = ([(n, C.POINTER(C.c_double)) for n in _F64] + [(n, C.POINTER(C.c_int64)) for n in _I64] +
                [(n, C.POINTER(C.c_int32)) for n in _I32] + [(n, C.POINTER(C.c_int32)) for n in _I32_TAIL] +
                [(n, C.POINTER(C.c_double)) for n in _F64_TAIL])


_lib = None


def load():
    """Loads libchs.so once; raises if it has not been built (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(the simulator has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, i32 = C.c_void_p, C.c_int32
    lib.chs_abi_version.restype = C.c_int
    lib.chs_last_error.restype = C.c_char_p
    lib.chs_create.argtypes = [C.POINTER(ChsParams), i32, i32, C.POINTER(vp)]
    lib.chs_destroy.argtypes = [vp]
    lib.chs_get_dims.argtypes = [vp, C.POINTER(ChsDims)]
    lib.chs_set_spawn_table.argtypes = [vp, vp, i32, i32]
    lib.chs_reset.argtypes = [vp, vp, vp, vp]
    lib.chs_step.argtypes = [vp] * 8
    lib.chs_step_static_scratch.argtypes = [vp] * 8
    lib.chs_step_host.argtypes = [vp] * 10
    lib.chs_reset_host.argtypes = [vp, vp, vp, vp]
    lib.chs_get_state.argtypes = [vp, i32, i32, C.POINTER(ChsStateView)]
    lib.chs_set_state.argtypes = [vp, i32, i32, C.POINTER(ChsStateView)]
    lib.chs_set_debug_buffers.argtypes = [vp, vp, vp]
    lib.chs_get_num_drones.argtypes = [vp, vp, vp]
    lib.chs_set_phase_clock.argtypes = [vp, vp]
    lib.chs_set_register_variant.argtypes = [vp, i32]
    lib.chs_set_step_mode.argtypes = [vp, i32]
    lib.chs_stats.argtypes = [vp, vp, i32]
    lib.chs_set_launch.argtypes = [vp, i32, i32]
    for s in SYMBOLS:
        if s not in ("chs_abi_version", "chs_last_error"):
            getattr(lib, s).restype = C.c_int
    if lib.chs_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libchs.so ABI {lib.chs_abi_version()} != binding ABI {ABI_VERSION}: rebuild")
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise RuntimeError(f"libchs error {rc}: {load().chs_last_error().decode()}")
