#!/usr/bin/env python
"""Benchmark of the batched RL-Cattle-Herding env step (BASELINE.json: env-steps/s, device timed).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs-per-gpu E] [--drones 3] [--cattle 20]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...      # the reference's own Python on the host cores (oracle/_ref)

A "step" is one control step of every environment of the batch = ONE launch of the fused CUDA kernel
(`chs_step`).  Workload at N=1: BASELINE.json configs[1], the 3-drone / 20-cattle scene (Physics.DYN, VEL
actions, 240 Hz physics / 30 Hz control, curriculum level 7, auto-reset) batched to 4096 envs per GPU; envs
are sharded by index over the ranks with no collective on the step path (weak scaling: 4096 envs per GPU).
Configs[2] (8 drones / 100 cattle x 1024 envs) and configs[3] (65 536 envs in total over the ranks) are measured in the
same process and reported under "configs".  Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--envs-per-gpu", type=int, default=4096)
    ap.add_argument("--drones", type=int, default=3)
    ap.add_argument("--cattle", type=int, default=20)
    ap.add_argument("--ctrl-freq", type=int, default=30)
    ap.add_argument("--launch", default="auto", help="'auto' or G,threads")
    ap.add_argument("--packing", default="ctde", choices=["ctde", "marl"],
                    help="ctde: sb3_envs CattleAviary (the BASELINE workload); marl: rllib_envs MARLCattleAviary semantics")
    ap.add_argument("--step-mode", default="auto", choices=["auto", "fused", "split"],
                    help="fused: one kernel per control step; split: drone_kernel + step_kernel; auto: the faster one (untimed probe)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip BASELINE configs 3 and 4")
    ap.add_argument("--no-parity-check", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the bounded CPU sample")
    return ap.parse_args()


def algorithmic_bytes_per_env_step(n, m, rows, obs_len, act_w=4):
    """fp32 bytes one env-step must move.  actions + drone state read/write (22 floats)
    + last RPM + herd read + herd velocity write every 2nd step + padded obs + reward/flags + env scalars."""
    return 4 * (n * act_w + 2 * n * 22 + n * 4 + m * 4 + 0.5 * m * 2 + rows * obs_len) + 4 + 2 + 2 * 16


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(workload_key):
    """dram bytes per launch of the step kernel from the committed ncu capture, if one matches the workload."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get(workload_key)
    return None


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ======================================================================================================
# CPU arm.  `oracle/ref_arm.py` times the reference's OWN code (the unmodified package from /root/reference, or its
# byte-for-byte staged copy oracle/_ref on the GPU box; pybullet is a stand-in) single-process and in
# a SubprocVecEnv-shaped pool; the float64 port (oracle/cattle_oracle.py) is timed beside it.  Checker / baseline use
# of oracle/ only: nothing below is on the product path.
# ======================================================================================================
def _ref_arm():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_arm
    return ref_arm


def _port_worker(args):
    seed, n, m, ctrl_freq, budget_s = args
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cattle_oracle as co
    import importlib.util
    spec = importlib.util.spec_from_file_location("chs_spawn", os.path.join(ROOT, "rl-cattle-herding_b200", "spawn.py"))
    spawn = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(spawn)
    table = spawn.generate(16 if m <= 32 else 4, m, seed=0)
    rng = np.random.default_rng(seed)
    env = co.OracleEnv(m, ctrl_freq=ctrl_freq, spawn_table=table, start_level=7)
    env.reset(n, rng.random((m, 2)))
    steps = 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < budget_s:
        a = rng.uniform(-1, 1, (n, 4)).astype(np.float32)
        _, _, term, trunc, _ = env.step(a)
        steps += 1
        if term or trunc:
            env.reset(n, rng.random((m, 2)))
    return steps, time.perf_counter() - t0


def port_env_steps_per_s(n, m, ctrl_freq, budget_s):
    steps, dt = _port_worker((1, n, m, ctrl_freq, budget_s))
    return steps / dt


def cpu_baseline_single(n, m, ctrl_freq, budget_s):
    """The `cpu_baseline` object of the GPU line: the reference, one process, on this box's host."""
    ra = _ref_arm()
    v, sample = ra.time_single(n, m, ctrl_freq, budget_s)
    out = {"value": v, "unit": "env-steps/s", "cores": 1, "kind": ra.kind(), "sample": sample}
    out.update(ra.host_info())
    out["third_party"] = ra.third_party()
    out["port"] = {"value": port_env_steps_per_s(n, m, ctrl_freq, min(4.0, budget_s)), "unit": "env-steps/s", "cores": 1,
                   "what": "oracle/cattle_oracle.py (float64 restatement), same workload"}
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ra = _ref_arm()
    host = ra.host_info()
    cores = host["cores_visible"]
    K, W = max(1, args.steps), args.warmup
    b = max(3.0, min(args.cpu_seconds, 20.0))
    n, m = args.drones, args.cattle
    single, single_s = ra.time_single(n, m, args.ctrl_freq, b)
    pool, used, pool_s = ra.time_pool(n, m, args.ctrl_freq, b * 1.5, cores)
    c3_single, c3_single_s = ra.time_single(8, 100, args.ctrl_freq, b)
    c3_pool, _, c3_pool_s = ra.time_pool(8, 100, args.ctrl_freq, b * 1.5, cores)
    port = port_env_steps_per_s(n, m, args.ctrl_freq, 4.0)
    line = {
        "metric": "env-steps/sec", "value": pool, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
        "ms_per_step": 1000.0 / pool, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "impl": "reference",
        "config": {"workload": f"CattleAviary Physics.DYN, {n} drones + {m} cattle, VEL actions, 240 Hz physics / "
                               f"{args.ctrl_freq} Hz ctrl, level 7, uniform random actions, auto-reset",
                   "note": "CPU arm: the reference's own unmodified Python (gym_pybullet_drones.sb3_envs.CattleAviary) on all "
                           "host cores, one env per process with one pipe round trip per step (SubprocVecEnv protocol, "
                           "simulator/CTDECattleHerder.py:91-97); pybullet is the stand-in body store of "
                           "oracle/ref_harness/standins (the real package is installed on neither box)"},
        "cpu_baseline": {"value": pool, "unit": "env-steps/s", "cores": used, "kind": ra.kind(), "sample": pool_s,
                         "single_process": {"value": single, "sample": single_s},
                         "c3_8d100c": {"single_process": {"value": c3_single, "sample": c3_single_s},
                                       "pool": {"value": c3_pool, "sample": c3_pool_s}},
                         "port": {"value": port, "cores": 1, "what": "oracle/cattle_oracle.py, single process, same workload"},
                         "third_party": ra.third_party(), **host},
        "e2e": {"value": pool, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ======================================================================================================
# GPU arm
# ======================================================================================================
L2_BYTES = 126e6          # B200 L2; a rotating set of batches is sized to 1.3x this


def time_isolated(env, torch, action_pool, n_steps, flush_buf):
    """Round-1 method, kept as a secondary figure: every launch timed on its own with a CUDA-event pair on the launching
    stream, L2 flushed between launches by zeroing a 256 MiB buffer outside the pairs.  An event pair around a single
    launch also times the launch itself (~4 us here)."""
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(n_steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(n_steps)]
    for i in range(n_steps):
        if flush_buf is not None:
            flush_buf.zero_()
        starts[i].record()
        env.step(action_pool[i % len(action_pool)])
        ends[i].record()
    torch.cuda.synchronize()
    return [s.elapsed_time(e) for s, e in zip(starts, ends)]   # ms each


class Rotating:
    """R independent batches of E envs of one scene, stepped round-robin on ONE stream.  R is chosen so that the bytes the
    R steps touch exceed L2 (state + live observation rows + actions per batch), i.e. every launch finds its inputs in
    HBM without an L2-flush kernel between launches ("inputs larger than L2"); consecutive launches are stream-ordered,
    never concurrent."""

    def __init__(self, pkg, torch, dev, E, N, M, ctrl_freq, rank_offset, packing="ctde", min_batches=1):
        self.torch, self.E, self.N, self.M = torch, E, N, M
        table = pkg.spawn.generate(100 if M <= 32 else 8, M, seed=0)
        self.table = table
        probe = dict(num_drones=N, num_cattle=M, ctrl_freq=ctrl_freq, fixed_num_drones=N, start_level=7, auto_reset=True,
                     spawn_table=table, device=dev.index, seed=0, packing=packing)
        first = pkg.CattleHerdingBatchedEnv(E, env_id_offset=rank_offset, **probe)
        d = first.dims
        self.dims = d
        rows_written = min(d["obs_rows"], d["max_drones"]) if first.persistent_padding else d["obs_rows"]
        self.touched_per_env = 4 * (N * 4 + 2 * N * 28 + N * 4 + M * 4 + M * 2 + rows_written * d["obs_len"]) + 70
        self.R = max(min_batches, int(np.ceil(1.3 * L2_BYTES / (self.touched_per_env * E))))
        self.envs = [first] + [pkg.CattleHerdingBatchedEnv(E, env_id_offset=rank_offset, **dict(probe, seed=r)) for r in range(1, self.R)]
        gen = torch.Generator(device=dev).manual_seed(1 + rank_offset)
        self.pool = [torch.rand((E, N, 4), device=dev, generator=gen) * 2 - 1 for _ in range(8 if E <= 65536 else 3)]
        for e in self.envs:
            e.reset()
        self.i = 0
        self.graph = None

    def step_once(self):
        self.envs[self.i % self.R].step(self.pool[self.i % len(self.pool)])
        self.i += 1

    def run(self, n):
        for _ in range(n):
            self.step_once()

    def configure(self, g, th, mode):
        for e in self.envs:
            e.set_step_mode(mode)
            e.set_launch(g, th)

    def time_stream(self, n):
        """n back-to-back `env.step` calls, one event pair; ms per step."""
        torch = self.torch
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        self.run(n)
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / n

    def time_graph(self, n):
        """n steps captured as one CUDA graph (the way the timed region runs them), replayed once untimed and once between
        an event pair; ms per step.  Shapes that differ by 1-2 % are not separable by eager launches from Python."""
        torch = self.torch
        g = self.capture(n)
        g.replay()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        self.graph = None
        return s.elapsed_time(e) / n

    def autotune(self, candidates, modes, reps=48):
        """Untimed: picks (envs per CTA, threads per CTA, step mode) by the time of a CUDA graph of `reps` steps; the three
        best are measured again, interleaved.  Returns (ms, G, threads, mode)."""
        reps = max(self.R, (reps // self.R) * self.R)          # whole rounds over the rotating batches
        ranked = []
        for mode in modes:
            for g, th in candidates:
                try:
                    self.configure(g, th, mode)
                    self.run(2 * self.R)
                    ranked.append((self.time_graph(reps), g, th, mode))
                except RuntimeError:
                    continue
        ranked.sort()
        acc = {f[1:]: [] for f in ranked[:3]}
        for _ in range(3):
            for key in acc:
                self.configure(*key)
                self.run(self.R)
                acc[key].append(self.time_graph(reps))
        best = min((float(np.median(v)), k[0], k[1], k[2]) for k, v in acc.items())
        self.configure(best[1], best[2], best[3])
        return best

    def capture(self, n):
        """One CUDA graph of n consecutive steps (n kernel nodes in fused mode, 2-3 n in split mode)."""
        torch = self.torch
        st = torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            self.run(self.R)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=st):
                self.run(n)
        torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()
        return self.graph

    def close(self):
        self.graph = None
        for e in self.envs:
            e.close()
        self.envs = []


def timed_region(rot, torch, dist, K, W, dev, use_graph=True):
    """W untimed warm-up steps, then EXACTLY K steps bracketed by barrier + synchronize, device-timed with one event pair
    on the launching stream, max over ranks.  Returns (ms_total_max_over_ranks, wall_s, clocks)."""
    graph = rot.capture(K) if use_graph else None
    rot.run(max(W, rot.R))
    if graph is not None:
        graph.replay()            # one untimed replay: the graph's own first-launch cost stays outside
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(dev.index) as clocks:
        t0 = time.perf_counter()
        s.record()
        if graph is not None:
            graph.replay()
        else:
            rot.run(K)
        e.record()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        wall = time.perf_counter() - t0
        if wall < 0.35:           # let nvidia-smi take a few samples while the same work repeats (untimed)
            t1 = time.perf_counter()
            while time.perf_counter() - t1 < 0.5:
                if graph is not None:
                    graph.replay()
                else:
                    rot.run(K)
                torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e)], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), wall, clocks.summary()


def cupti_kernel_ms(rot, torch, n=24):
    """The same steps as CUPTI's kernel activity records see them (sum of our kernels per step), untimed, supplementary."""
    try:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            rot.run(n)
            torch.cuda.synchronize()
        ours = [ev.time_range.elapsed_us() for ev in prof.events()
                if ev.device_type == torch.autograd.DeviceType.CUDA and "chs::" in ev.name]
        return sum(ours) / n * 1e-3 if ours else None
    except Exception:  # noqa: BLE001  (profiler not available: the field stays null)
        return None


def roofline_block(N, M, dims, E, ms, mode, peak, peak_src, traffic_key):
    padded = algorithmic_bytes_per_env_step(N, M, dims["obs_rows"], dims["obs_len"])
    compact = algorithmic_bytes_per_env_step(N, M, min(dims["obs_rows"], dims["max_drones"]), dims["obs_len"])
    ach = padded * E / (ms * 1e-3) / 1e9
    ach_c = compact * E / (ms * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": recorded_traffic(traffic_key), "bytes_per_env_step": padded, "bytes_per_launch": padded * E,
            "accounting": "SURVEY 8(d) padded-observation figure (the (12, L) image of the reference); with persistent_padding "
                          "the constant rows >= max_drones are written at reset only -- `compact` is the same time against the "
                          "bytes a step actually has to move",
            "compact": {"bytes_per_env_step": compact, "achieved": ach_c, "frac": ach_c / peak},
            "peak_source": peak_src, "kernel_ms": ms,
            "kernel": "chs::step_kernel<float> (one launch per control step)" if mode == "fused" else
                      "chs::drone_kernel<float> + chs::step_kernel<float, SPLIT>: one control step"}


def launches_per_step(mode):
    return 1 if mode == "fused" else 2


def measure_config(pkg, torch, dist, dev, rank, E, N, M, ctrl_freq, K, W, cands, modes, peak, peak_src, key):
    """One extra configuration measured like the headline (rotating batches, graph of K steps, max over ranks)."""
    rot = Rotating(pkg, torch, dev, E, N, M, ctrl_freq, rank * E)
    tuned = rot.autotune(cands, modes, reps=max(12, min(48, 2 * rot.R)))
    total_ms, _, _ = timed_region(rot, torch, dist, K, W, dev)
    ms = total_ms / K
    world = dist.get_world_size() if dist is not None else 1
    out = {"envs_per_gpu": E, "total_envs": E * world, "drones": N, "cattle": M, "ms_per_step": ms,
           "env_steps_per_s": world * E / (ms * 1e-3), "batches_rotated": rot.R,
           "launch": {"envs_per_block": tuned[1], "threads_per_block": tuned[2], "step_mode": tuned[3]},
           "roofline": roofline_block(N, M, rot.dims, E, ms, tuned[3], peak, peak_src, key)}
    out["roofline"]["kernel_activity_ms_cupti"] = cupti_kernel_ms(rot, torch)
    rot.close()
    return out


def pin_rank_to_cores(local_rank, world):
    """One process per GPU on one socket: give each rank its own slice of the visible cores so that the launch threads
    of the ranks do not migrate onto each other."""
    try:
        cores = sorted(os.sched_getaffinity(0))
        k = max(1, len(cores) // max(1, world))
        mine = cores[local_rank * k:(local_rank + 1) * k] or cores
        os.sched_setaffinity(0, mine)
        return mine
    except (AttributeError, OSError):
        return None


def run_cuda_arm(args):
    import torch
    import rl_cattle_herding_b200 as pkg
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    cores = pin_rank_to_cores(local_rank, world) if world > 1 else None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        torch.cuda.set_device(0)
    dev = torch.device("cuda", torch.cuda.current_device())
    E, N, M = args.envs_per_gpu, args.drones, args.cattle
    K, W = args.steps, max(3, args.warmup)
    peak, peak_src = measured_peak_gbs()

    # ---- headline: BASELINE configs[1] ----
    rot = Rotating(pkg, torch, dev, E, N, M, args.ctrl_freq, rank * E, packing=args.packing)
    if args.launch == "auto":
        # (28, T): wave-fitted shapes -- a CTA handles 28 of its 32 env slots, so that 4096 envs make 147 CTAs for the 148 SMs
        # (one CTA of 512 threads per SM) and 16 384 envs make 586 CTAs for the 592 resident slots of the 64-register build
        cands = ([(8, 128), (16, 256), (8, 96), (4, 64), (4, 128), (16, 128), (32, 512), (28, 512), (28, 256)] if E <= 16384
                 else [(32, 256), (16, 256), (16, 128)])
        modes = ("fused", "split") if args.step_mode == "auto" and E > 16384 else (("fused",) if args.step_mode == "auto" else (args.step_mode,))
        tuned = rot.autotune(cands, modes)
    else:
        g, th = (int(x) for x in args.launch.split(","))
        mode = "fused" if args.step_mode == "auto" else args.step_mode
        rot.configure(g, th, mode)
        tuned = (None, g, th, mode)
    for e in rot.envs:
        e.stats(reset=True)
    total_ms, t_wall, clocks = timed_region(rot, torch, dist, K, W, dev)
    ms_per_step = total_ms / K
    value = world * E * K / (total_ms * 1e-3)
    stats = torch.from_numpy(np.sum([e.stats() for e in rot.envs], axis=0)).to(dev)
    if dist is not None:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)           # the per-iteration episode-statistics reduction (NCCL)

    # secondary figures for the same launches (untimed region): plain stream launches, isolated launches with an L2 flush
    stream_ms = rot.time_stream(max(48, 2 * rot.R))
    flush_buf = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    iso = time_isolated(rot.envs[0], torch, rot.pool, 40, flush_buf)
    iso_ms = float(np.mean(sorted(iso)[:30]))
    del flush_buf
    act = torch.tensor([cupti_kernel_ms(rot, torch) or 0.0], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(act, op=dist.ReduceOp.MAX)
    activity_ms = float(act.item()) or None

    line = None
    dims = rot.dims
    if rank == 0:
        wl_key = f"{N}d{M}c_{E}env"
        line = {
            "metric": "env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"CattleAviary Physics.DYN, {N} drones + {M} cattle, {E} envs per GPU, VEL actions, "
                                   f"240 Hz physics / {args.ctrl_freq} Hz ctrl, level 7, uniform random actions, auto-reset"
                                   + (", MARL packing" if args.packing == "marl" else ""),
                       "envs_per_gpu": E, "total_envs": world * E, "obs_shape": [dims["obs_rows"], dims["obs_len"]],
                       "launch": {"envs_per_block": tuned[1], "threads_per_block": tuned[2], "smem_bytes": rot.envs[0].dims["smem_bytes"],
                                  "step_mode": tuned[3]},
                       "persistent_padding": bool(rot.envs[0].persistent_padding),
                       "l2": f"inputs larger than L2: {rot.R} independent {E}-env batches stepped round-robin on one stream "
                             f"({rot.R * rot.touched_per_env * E / 1e6:.0f} MB touched per round > 126 MB L2), no flush kernel",
                       "timing": "one CUDA-event pair on the launching stream around a CUDA graph of exactly K consecutive steps "
                                 "(one kernel node per step), barrier + synchronize on both sides, max over ranks",
                       "launch_mode": ("plain launches (CHS_NO_PDL)" if os.environ.get("CHS_NO_PDL") else
                                       "programmatic dependent launch: a step kernel's launch and shared-memory prologue overlap the "
                                       "tail of the previous kernel of the stream; it executes griddepcontrol.wait (the previous "
                                       "kernel complete, its memory visible) before its first global-memory access"),
                       "sharding": "env index, no collective on the step path; one NCCL all-reduce of the stats vector after the run",
                       "rank_cores": cores},
            "drone_steps_per_s": value * N,
            "wall_ms_per_step": t_wall * 1e3 / K,
            "ms_per_step_stream_launches": stream_ms,
            "ms_per_step_isolated_l2_flushed": iso_ms,
            "gpu_launches": K * launches_per_step(tuned[3]),
            "roofline": roofline_block(N, M, dims, E, ms_per_step, tuned[3], peak, peak_src, wl_key),
            "clocks": clocks,
            "episodes": {"finished": float(stats[0].item()), "mean_length": float(stats[2].item() / max(1.0, stats[0].item())),
                         "mean_return": float(stats[1].item() / max(1.0, stats[0].item())), "non_finite": float(stats[9].item())},
        }
        line["roofline"]["kernel_activity_ms_cupti"] = activity_ms
        line["roofline"]["isolated"] = {"kernel_ms": iso_ms, "frac": algorithmic_bytes_per_env_step(N, M, dims["obs_rows"], dims["obs_len"])
                                        * E / (iso_ms * 1e-3) / 1e9 / peak, "what": "round-1 method: event pair per launch, L2 flushed"}

    # ---- parity stamp of the timed configuration (untimed; oracle = checker only) ----
    if rank == 0 and args.packing == "ctde" and not args.no_parity_check:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import subset_check as sc
        env0 = rot.envs[0]
        chk = sc.SubsetChecker(env0, rot.table, sc.subset_ids(E, 16, 4), seed=0, env_id_offset=0, fixed_num_drones=N)
        for i in range(4):
            chk.step(rot.pool[i % len(rot.pool)])
        line["parity_check"] = dict(chk.summary(), against="oracle/cattle_oracle.py (pinned to the reference's recordings) stepped from "
                                    "chs_get_state() of the bench's own env; tolerances of tests/test_cuda_parity.py")
        line["parity_check"]["failures"] = [str(f) for f in chk.failures[:3]]

    # ---- e2e: the numpy-facing call (pinned host actions in, pinned host obs/reward/flags out), per rank ----
    if not args.no_e2e:
        env0 = rot.envs[0]
        host_actions = [a.cpu().numpy() for a in rot.pool[:4]]
        env0.reset_host()
        for i in range(5):
            env0.step_host(host_actions[i % 4])
        np.copyto(env0.host_actions, host_actions[0])
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        Ke = max(10, min(K, 200))
        # five repetitions of the same Ke-step region, the median one is reported (a 20 ms wall-clock region on a shared host
        # is exposed to scheduler noise: single regions of one build ranged from 3.5e7 to 4.3e7 env-steps/s between boxes)
        reps, nterm = [], 0
        for _ in range(5):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for i in range(Ke):
                out = env0.step_host()      # actions already sit in the pinned buffer (e2e contract: inputs start in pinned memory)
                nterm += len(out[4])
            torch.cuda.synchronize()
            reps.append(time.perf_counter() - t0)
        nterm //= 5
        e2e_s = torch.tensor([float(np.median(reps))], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
        if rank == 0:
            D = dims["max_drones"]
            live = (min(dims["obs_rows"], D) * dims["obs_len"] * 4 + 15) // 16 * 16      # live rows of one image, as stored
            small = 16 + E * 4 + 2 * ((E + 15) // 16 * 16) + E * 4
            line["e2e"] = {"value": world * E * Ke / float(e2e_s.item()), "unit": "env-steps/s",
                           "h2d_bytes_per_step": E * N * 4 * 4,
                           "d2h_bytes_per_step": E * live + small + int(nterm / Ke * live),
                           "steps": Ke, "repetitions_us_per_step": [round(1e6 * r / Ke, 2) for r in reps], "per_rank_d2h_gbs": (E * live + small + nterm / Ke * live) * Ke / float(e2e_s.item()) / 1e9,
                           "api": "CattleHerdingBatchedEnv.step_host (chs_step_host): the kernel reads the actions from, and writes "
                                  "observations / terminal observations / reward / flags / done indices straight into, the caller's pinned "
                                  "buffers (no staging, no copy node: one launch and one stream synchronisation per step)"}
    rot.close()

    # ---- BASELINE configs[2] and configs[3], same method, every rank ----
    if not args.no_configs:
        c3 = measure_config(pkg, torch, dist, dev, rank, 1024, 8, 100, args.ctrl_freq, K, W,
                            [(4, 256), (2, 128), (2, 256), (4, 128)], ("fused",), peak, peak_src, "8d100c_1024env")
        Ec4 = 65536 // world
        c4 = measure_config(pkg, torch, dist, dev, rank, Ec4, 3, 20, args.ctrl_freq, min(K, 100), W,
                            [(32, 256), (16, 256)] if Ec4 > 16384 else [(16, 256), (8, 128), (16, 128), (28, 256), (32, 256), (28, 512)],
                            ("fused", "split") if Ec4 >= 16384 else ("fused",), peak, peak_src, f"3d20c_{Ec4}env")
        c4["scaling"] = "strong: 65 536 envs in total, split evenly by env index"
        if rank == 0:
            line["configs"] = {"c3_8d100c_1024env": c3, "c4_65536_total": c4}

    # ---- larger batches on the same GPU, rank 0 of N=1 only ----
    if rank == 0 and world == 1 and not args.no_sweep:
        sweep = []
        for Es in (16384, 262144):
            if Es == E:
                continue
            r = measure_config(pkg, torch, None, dev, 0, Es, N, M, args.ctrl_freq, 60, W,
                               [(16, 256), (32, 256), (28, 256), (16, 128)], ("fused", "split"), peak, peak_src, f"{N}d{M}c_{Es}env")
            sweep.append({"envs": Es, "ms_per_step": r["ms_per_step"], "env_steps_per_s": r["env_steps_per_s"],
                          "hbm_frac": r["roofline"]["frac"], "hbm_frac_compact": r["roofline"]["compact"]["frac"],
                          "launch": r["launch"], "batches_rotated": r["batches_rotated"]})
        if "configs" in line:
            c4 = line["configs"]["c4_65536_total"]
            sweep.append({"envs": 65536, "ms_per_step": c4["ms_per_step"], "env_steps_per_s": c4["env_steps_per_s"],
                          "hbm_frac": c4["roofline"]["frac"], "hbm_frac_compact": c4["roofline"]["compact"]["frac"],
                          "launch": c4["launch"], "batches_rotated": c4["batches_rotated"]})
        line["roofline"]["sweep"] = sorted(sweep, key=lambda r: r["envs"])

    # ---- CPU baseline beside it (rank 0, N=1): the reference's own code on this box's host ----
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_single(N, M, args.ctrl_freq, args.cpu_seconds)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_cuda_arm(args)


if __name__ == "__main__":
    main()
