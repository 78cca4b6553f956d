"""`CattleHerdingBatchedEnv`: E independent CattleAviary environments stepped by one CUDA launch.

Thin host layer over the C ABI (`include/chs.h`).  PyTorch is used for device memory, streams and
(in `dist.py`) `torch.distributed`; all simulation arithmetic happens inside libchs.so.  Semantics are those
of the reference's `CattleAviary.step()/reset()` on its `Physics.DYN` path
(gym_pybullet_drones/sb3_envs/BaseAviary.py:280-465, BaseRLAviary.py:137-342, CattleAviary.py:213-552).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .enums import ACT_CODE, ACT_WIDTH

STATE_F64 = _lib._F64 + _lib._F64_TAIL
STATE_INT = _lib._I64 + _lib._I32 + _lib._I32_TAIL


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class CattleHerdingBatchedEnv:
    """Batched DYN-path simulator.

    Parameters mirror `CattleAviary.__init__` (sb3_envs/CattleAviary.py:14-28): `num_drones` sizes the action
    space (drone slots), `num_cattle`, `pyb_freq`, `ctrl_freq`, `act`.  Additional knobs: `num_envs`,
    `start_level` (the reference hard-codes 7), `fixed_num_drones` (0 = draw per episode in the level's
    [min,max] like BaseAviary.py:307), `obs_rows` (12 = reference padding), `auto_reset`, `precision`
    ("fp32" product / "fp64" parity instrument), `env_id_offset` (global id of env 0, for sharding), `seed`.

    Non-reference options, both off by default: `aero` (bit mask or names "drag", "ground_effect", "downwash":
    the PYB-only aerodynamic models of sb3_envs/BaseAviary.py:943-1039 as forces of the DYN step) and
    `cattle_integrate` (herd positions advance by vel * CTRL_TIMESTEP every control step).

    `packing="marl"` selects the rllib_envs semantics (MARLCattleAviary / BaseMARLAviary): per-drone
    observation rows without the 12-slot neighbour gap, `reward` (E, D), `terminated` / `truncated` (E, D+1)
    whose last column is RLlib's "__all__" entry; `auto_reset` fires on either "__all__" flag.

    `persistent_padding` (default on): rows >= num_drones of the (12, L) observation images are constant zeros
    (`BaseRLAviary._computeObs` pads every observation to GLOBAL_MAX_NUM_DRONES rows, BaseRLAviary.py:334).  The
    tensors `self.obs` / `self.terminal_obs` (and the pinned host mirrors) are owned by this object and are created
    zero-filled, so the simulator writes those rows at `reset()` only and never again: 3/4 of the bytes of an image
    at 3 drones.  Turn it off to have every step rewrite the full image.

    `marl_wrapper_quirks` (MARL packing, default off): reward / flags are what the reference's RLlib wrapper reports --
    the second invocation of the stateful per-agent functions (rllib_envs/marl_wrapper.py:104-117), see include/chs.h.
    """

    def __init__(self, num_envs, num_drones=3, num_cattle=20, pyb_freq=240, ctrl_freq=30, act="vel",
                 start_level=7, fixed_num_drones=0, obs_rows=None, auto_reset=True, precision="fp32",
                 spawn_table=None, device=None, env_id_offset=0, seed=0, envs_per_block=0, threads_per_block=0,
                 packing="ctde", aero=0, cattle_integrate=False, step_mode="auto", persistent_padding=True,
                 marl_wrapper_quirks=False):
        if not torch.cuda.is_available():
            raise RuntimeError("CattleHerdingBatchedEnv needs a CUDA device (there is no CPU fallback)")
        self._ctor = dict(num_envs=num_envs, num_drones=num_drones, num_cattle=num_cattle, pyb_freq=pyb_freq, ctrl_freq=ctrl_freq,
                          act=act, start_level=start_level, fixed_num_drones=fixed_num_drones, obs_rows=obs_rows,
                          auto_reset=auto_reset, precision=precision, device=device, env_id_offset=env_id_offset, seed=seed,
                          packing=packing, aero=aero, cattle_integrate=cattle_integrate, step_mode=step_mode,
                          persistent_padding=persistent_padding, marl_wrapper_quirks=marl_wrapper_quirks)
        self._spawn_table = None
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else
                                   (device if isinstance(device, int) else torch.device(device).index or 0))
        act_code = ACT_CODE[act]
        p = _lib.ChsParams()
        p.struct_size = C.sizeof(_lib.ChsParams)
        p.num_cattle, p.max_drones, p.fixed_num_drones = int(num_cattle), int(num_drones), int(fixed_num_drones)
        p.pyb_freq, p.ctrl_freq, p.act_type, p.start_level = int(pyb_freq), int(ctrl_freq), act_code, int(start_level)
        self.packing = packing
        p.step_mode = {"auto": 0, "fused": 1, "split": 2}[step_mode]
        if not isinstance(aero, int):       # names -> ChsParams.aero_flags bits
            aero = sum({"drag": 1, "ground_effect": 2, "downwash": 4}[str(x)] for x in aero)
        p.aero_flags, p.cattle_integrate = int(aero), int(bool(cattle_integrate))
        self.aero_flags, self.cattle_integrate = int(aero), bool(cattle_integrate)
        if obs_rows is None:
            obs_rows = 12 if packing == "ctde" else int(num_drones)
        p.obs_rows, p.packing, p.precision = int(obs_rows), {"ctde": 0, "marl": 1}[packing], {"fp32": 0, "fp64": 1}[precision]
        p.auto_reset, p.env_id_offset, p.seed = int(bool(auto_reset)), int(env_id_offset), int(seed)
        p.envs_per_block, p.threads_per_block = int(envs_per_block), int(threads_per_block)
        p.persistent_padding = int(bool(persistent_padding))
        p.marl_wrapper_quirks = int(bool(marl_wrapper_quirks))
        self.persistent_padding = bool(persistent_padding)
        self._h = C.c_void_p(0)
        _lib.check(self.lib.chs_create(C.byref(p), int(num_envs), self.device.index, C.byref(self._h)))
        d = _lib.ChsDims()
        _lib.check(self.lib.chs_get_dims(self._h, C.byref(d)))
        self.num_envs, self.max_drones, self.num_cattle = d.num_envs, d.max_drones, d.num_cattle
        self.action_width, self.obs_rows, self.obs_len, self.substeps = d.action_width, d.obs_rows, d.obs_len, d.substeps
        assert self.action_width == ACT_WIDTH[act_code]
        self.reward_width, self.flag_width = d.reward_width, d.flag_width
        self.auto_reset = bool(auto_reset)
        self.precision = precision
        self.ctrl_freq, self.pyb_freq = int(ctrl_freq), int(pyb_freq)
        E, dev = self.num_envs, self.device
        # tensors owned by the host layer; the kernel writes straight into them
        self.obs = torch.zeros((E, self.obs_rows, self.obs_len), dtype=torch.float32, device=dev)
        self.terminal_obs = torch.zeros_like(self.obs)
        rshape = (E,) if packing == "ctde" else (E, self.reward_width)
        fshape = (E,) if packing == "ctde" else (E, self.flag_width)
        self._rshape, self._fshape = rshape, fshape
        self.reward = torch.zeros(rshape, dtype=torch.float32, device=dev)
        self.terminated = torch.zeros(fshape, dtype=torch.uint8, device=dev)
        self.truncated = torch.zeros(fshape, dtype=torch.uint8, device=dev)
        self._dbg_nb = self._dbg_herd = None
        self._host = None
        if spawn_table is not None:
            self.set_spawn_table(spawn_table)

    # ------------------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            torch.cuda.synchronize(self.device)
            self.lib.chs_destroy(self._h)
            self._h = C.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def dims(self):
        d = _lib.ChsDims()
        _lib.check(self.lib.chs_get_dims(self._h, C.byref(d)))
        return {n: getattr(d, n) for n, _ in d._fields_}

    def set_step_mode(self, mode):
        """"fused": one kernel per control step; "split": drone_kernel + step_kernel; "auto": by batch size."""
        _lib.check(self.lib.chs_set_step_mode(self._h, {"auto": 0, "fused": 1, "split": 2}[mode]))

    def set_launch(self, envs_per_block, threads_per_block):
        _lib.check(self.lib.chs_set_launch(self._h, int(envs_per_block), int(threads_per_block)))

    # launch shapes `autotune_launch` tries by default: (envs per CTA, threads per CTA).  28 = wave fitting (include/chs.h,
    # chs_set_launch): 147 CTAs for the 148 SMs at 4096 envs, 586 for the 592 resident slots at 16 384
    LAUNCH_CANDIDATES = ((2, 128), (4, 64), (4, 128), (4, 256), (8, 128), (8, 256), (16, 128), (16, 256), (28, 256), (32, 256),
                         (28, 512), (32, 512))

    def autotune_launch(self, candidates=None, modes=None, steps=24):
        """Picks the launch shape (and, when `modes` names several, the step mode) that steps a batch like this one fastest
        on this GPU, and applies it to this env.  The timing runs on a PROBE env built from the same constructor arguments
        (same batch size, spawn table and seeds; destroyed afterwards), so this env's state, episode counters and statistics
        are untouched: every candidate steps the probe under uniform random actions, `steps` steps as one CUDA graph between
        CUDA events, and the two best are measured again, interleaved.  Results do not depend on the shape
        (tests/test_cuda_env.py), only the time does -- the best shape depends on how the batch fills whole waves of the
        148 SMs.  Returns (ms per step, envs per CTA, threads per CTA, mode)."""
        if self._spawn_table is None:
            raise RuntimeError("set_spawn_table() first: the probe steps real episodes")
        cands = list(candidates or self.LAUNCH_CANDIDATES)
        modes = list(modes or [self._ctor["step_mode"]])
        probe = CattleHerdingBatchedEnv(spawn_table=self._spawn_table, **self._ctor)
        try:
            act = [torch.rand((self.num_envs, self.max_drones, self.action_width), device=self.device) * 2 - 1 for _ in range(4)]
            probe.reset()

            def timed(g, th, mode, n):
                # the n steps run as ONE CUDA graph: eager launches from Python cannot separate shapes that differ by 1-2 %
                probe.set_launch(g, th)
                probe.set_step_mode(mode)
                side = torch.cuda.Stream(device=self.device)
                side.wait_stream(torch.cuda.current_stream(self.device))
                with torch.cuda.stream(side):
                    for i in range(8):
                        probe.step(act[i % 4], want_terminal_obs=False)
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, stream=side):
                        for i in range(n):
                            probe.step(act[i % 4], want_terminal_obs=False)
                torch.cuda.current_stream(self.device).wait_stream(side)
                graph.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                graph.replay()
                e1.record()
                e1.synchronize()
                return e0.elapsed_time(e1) / n

            ranked = []
            for mode in modes:
                for g, th in cands:
                    try:
                        ranked.append((timed(g, th, mode, steps), g, th, mode))
                    except RuntimeError:        # shape refused for this batch (shared memory, occupancy variant)
                        continue
            if not ranked:
                raise RuntimeError("no launch candidate was accepted")
            ranked.sort()
            again = {r[1:]: [] for r in ranked[:2]}
            for _ in range(3):
                for key in again:
                    again[key].append(timed(*key, steps))
            best = min((float(np.median(v)),) + k for k, v in again.items())
        finally:
            probe.close()
        self.set_launch(best[1], best[2])
        self.set_step_mode(best[3])
        return best

    def set_spawn_table(self, table):
        t = np.ascontiguousarray(np.asarray(table, dtype=np.float64))
        assert t.ndim == 3 and t.shape[2] == 2, "spawn table must be (n_sims, cows_per_sim, 2)"
        self._spawn_table = t
        _lib.check(self.lib.chs_set_spawn_table(self._h, t.ctypes.data_as(C.c_void_p), t.shape[0], t.shape[1]))

    def enable_debug_outputs(self):
        """Integer diagnostics compared bit-exactly by the parity tests (neighbour ranks, herded count)."""
        self._dbg_nb = torch.full((self.num_envs, self.max_drones, 2), -1, dtype=torch.int32, device=self.device)
        self._dbg_herd = torch.zeros((self.num_envs,), dtype=torch.int32, device=self.device)
        _lib.check(self.lib.chs_set_debug_buffers(self._h, _ptr(self._dbg_nb), _ptr(self._dbg_herd)))
        return self._dbg_nb, self._dbg_herd

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # ------------------------------------------------------------------------------------------------
    def reset(self, mask=None):
        """`BaseAviary.reset` for the masked envs (all when None).  Returns the obs tensor (E, R, L)."""
        if mask is not None:
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
        _lib.check(self.lib.chs_reset(self._h, _ptr(mask), _ptr(self.obs), self._stream()))
        return self.obs

    def step(self, actions, want_terminal_obs=True):
        """One control step of every env.  `actions`: float32 CUDA tensor (E, D, A).
        Returns (obs, reward, terminated, truncated) -- views of tensors owned by this object; with
        `auto_reset` finished envs already hold the first observation of their next episode and
        `self.terminal_obs[i]` holds the last one of the finished episode."""
        if actions.dtype != torch.float32 or not actions.is_cuda or not actions.is_contiguous():
            actions = actions.to(device=self.device, dtype=torch.float32).contiguous()
        assert actions.numel() == self.num_envs * self.max_drones * self.action_width, "actions must be (E, D, A)"
        tobs = self.terminal_obs if (want_terminal_obs and self.auto_reset) else None
        _lib.check(self.lib.chs_step(self._h, _ptr(actions), _ptr(self.obs), _ptr(self.reward), _ptr(self.terminated),
                                     _ptr(self.truncated), _ptr(tobs), self._stream()))
        return self.obs, self.reward, self.terminated, self.truncated

    def num_drones_dev(self):
        """`NUM_DRONES` of every env as an int32 CUDA tensor (E,) owned by this object (stream-ordered copy of the
        simulator's per-env scalar; BaseAviary.py:307 draws it per episode)."""
        if getattr(self, "_ndr", None) is None:
            self._ndr = torch.zeros((self.num_envs,), dtype=torch.int32, device=self.device)
        _lib.check(self.lib.chs_get_num_drones(self._h, _ptr(self._ndr), self._stream()))
        return self._ndr

    def step_static_scratch(self, actions, want_terminal_obs=True):
        """`step()` through `chs_step_static_scratch`: the split-mode kernels with their scratch in the library's static
        device array (8 MiB per GPU, shared by every env object of the process), for callers that must not allocate.
        Raises if the batch does not fit; calls that overlap on the GPU race on the array."""
        if actions.dtype != torch.float32 or not actions.is_cuda or not actions.is_contiguous():
            actions = actions.to(device=self.device, dtype=torch.float32).contiguous()
        assert actions.numel() == self.num_envs * self.max_drones * self.action_width, "actions must be (E, D, A)"
        tobs = self.terminal_obs if (want_terminal_obs and self.auto_reset) else None
        _lib.check(self.lib.chs_step_static_scratch(self._h, _ptr(actions), _ptr(self.obs), _ptr(self.reward),
                                                    _ptr(self.terminated), _ptr(self.truncated), _ptr(tobs), self._stream()))
        return self.obs, self.reward, self.terminated, self.truncated

    # ---- host-buffer path (numpy in / numpy out; what a VecEnv calls) --------------------------------
    def _host_buffers(self):
        if self._host is None:
            E, R, L = self.num_envs, self.obs_rows, self.obs_len
            pin = dict(pin_memory=True)
            self._host = {
                "actions": torch.zeros((E, self.max_drones, self.action_width), dtype=torch.float32, **pin),
                "obs": torch.zeros((E, R, L), dtype=torch.float32, **pin),
                "reward": torch.zeros(self._rshape, dtype=torch.float32, **pin),
                "terminated": torch.zeros(self._fshape, dtype=torch.uint8, **pin),
                "truncated": torch.zeros(self._fshape, dtype=torch.uint8, **pin),
                "done_idx": torch.zeros((E,), dtype=torch.int32, **pin),
                "terminal_obs": torch.zeros((E, R, L), dtype=torch.float32, **pin),
            }
            self._host_np = {k: v.numpy() for k, v in self._host.items()}
            self._n_done = C.c_int32(0)
            # the argument list of chs_step_host never changes (the buffers are owned by this object): built once, so that a
            # call costs one ctypes dispatch instead of ten pointer conversions
            h = self._host
            self._host_args = (self._h, _ptr(h["actions"]), _ptr(h["obs"]), _ptr(h["reward"]), _ptr(h["terminated"]),
                               _ptr(h["truncated"]), C.byref(self._n_done), _ptr(h["done_idx"]), _ptr(h["terminal_obs"]))
        return self._host_np

    def reset_host(self, mask=None):
        hb = self._host_buffers()
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
        _lib.check(self.lib.chs_reset_host(self._h, C.c_void_p(0) if m is None else m.ctypes.data_as(C.c_void_p),
                                           C.c_void_p(self._host["obs"].data_ptr()), self._stream()))
        return hb["obs"]

    @property
    def host_actions(self):
        """The pinned (E, D, A) float32 action buffer of the host path.  Fill it in place and call
        `step_host()` without an argument (or with this very array) to skip the copy of the actions."""
        return self._host_buffers()["actions"]

    def step_host(self, actions=None):
        """numpy actions (E, D, A) -> numpy (obs, reward, terminated, truncated, done_idx, terminal_obs[:n]).
        The returned arrays are views of PINNED buffers owned by this object, overwritten by the next call.  The
        kernel reads the actions from, and writes the observations and terminal observations straight into, those
        pinned buffers over PCIe (no staging in HBM, no separate copy); reward / flags / done indices come back in one
        small copy.  `actions=None` (or the `host_actions` array itself): the pinned action buffer was filled in place."""
        hb = self._host_buffers()
        if actions is not None and actions is not hb["actions"]:
            np.copyto(hb["actions"], np.asarray(actions, dtype=np.float32).reshape(hb["actions"].shape))
        rc = self.lib.chs_step_host(*self._host_args, self._stream())
        if rc:
            _lib.check(rc)
        n = self._n_done.value
        return hb["obs"], hb["reward"], hb["terminated"], hb["truncated"], hb["done_idx"][:n], hb["terminal_obs"][:n]

    # ---- state I/O ------------------------------------------------------------------------------------
    def _view(self, arrays):
        v = _lib.ChsStateView()
        for k, a in arrays.items():
            ct = C.c_double if k in STATE_F64 else (C.c_int64 if k in _lib._I64 else C.c_int32)
            setattr(v, k, a.ctypes.data_as(C.POINTER(ct)))
        return v

    def get_state(self, env0=0, count=None):
        """Full simulator state of envs [env0, env0+count) as float64/int numpy arrays (env-major), the fields
        of `oracle/ref_harness/ref_env.capture_state`."""
        n = self.num_envs - env0 if count is None else count
        D, M = self.max_drones, self.num_cattle
        shp = {"pos": (n, D, 3), "quat": (n, D, 4), "rpy": (n, D, 3), "vel": (n, D, 3), "ang_v": (n, D, 3),
               "rpy_rates": (n, D, 3), "last_rpm": (n, D, 4), "pid_ipos": (n, D, 3), "pid_last_rpy": (n, D, 3),
               "pid_irpy": (n, D, 3), "cattle_pos": (n, M, 3), "cattle_vel": (n, M, 3), "prev_cent_dist": (n,),
               "spacing_clock": (n,), "ep_return": (n,), "path_len": (n, D)}
        out = {k: np.zeros(s, dtype=np.float64) for k, s in shp.items()}
        out["step_counter"] = np.zeros((n,), dtype=np.int64)
        for k in _lib._I32 + _lib._I32_TAIL:
            out[k] = np.zeros((n,), dtype=np.int32)
        v = self._view(out)
        _lib.check(self.lib.chs_get_state(self._h, int(env0), int(n), C.byref(v)))
        return out

    def set_state(self, state, env0=0):
        """Injects (a subset of) the state fields for envs [env0, env0+n); n is taken from the arrays."""
        n = None
        arrays = {}
        for k, a in state.items():
            if k not in STATE_F64 and k not in STATE_INT:
                continue
            dt = np.float64 if k in STATE_F64 else (np.int64 if k in _lib._I64 else np.int32)
            a = np.ascontiguousarray(np.asarray(a), dtype=dt)
            arrays[k] = a
            n = a.shape[0] if n is None else n
            assert a.shape[0] == n, "all state arrays need the same leading (env) dimension"
        v = self._view(arrays)
        _lib.check(self.lib.chs_set_state(self._h, int(env0), int(n), C.byref(v)))

    def stats(self, reset=False):
        out = np.zeros((_lib.NUM_STATS,), dtype=np.float64)
        _lib.check(self.lib.chs_stats(self._h, out.ctypes.data_as(C.c_void_p), int(bool(reset))))
        return out
