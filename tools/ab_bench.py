"""A/B timing of library builds on the GPU box: step time of the bench's shapes, measured the way bench.py measures
them (rotating batches larger than L2, one CUDA graph of K steps, one event pair), for the library named by CHS_LIB.
    CHS_LIB=/path/libchs_variant.so python tools/ab_bench.py [head] [c3] [mid] [c4] [big]
Launch shapes: the ones bench.py's autotune settles on (fixed here so that variants are compared like for like) unless
AB_SHAPES="G,T;G,T" overrides the launch shapes of every case named, AB_MODE=fused|split the step mode, AB_ENVS the batch size, AB_VARIANT=1|2 the register build."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import rl_cattle_herding_b200 as pkg  # noqa: E402

what = [a for a in sys.argv[1:]] or ["head", "c3", "mid", "c4"]
dev = torch.device("cuda", 0)
K = int(os.environ.get("AB_STEPS", "200"))
CASES = {"head": (4096, 3, 20, [(16, 256), (8, 128)], "fused"), "c3": (1024, 8, 100, [(4, 256), (8, 512)], "fused"),
         "mid": (16384, 3, 20, [(32, 256)], "fused"), "c4": (65536, 3, 20, [(32, 256)], "fused"),
         "big": (262144, 3, 20, [(32, 256)], "split")}
tag = os.path.basename(os.environ.get("CHS_LIB", "libchs.so"))
for name in what:
    E, N, M, shapes, mode = CASES[name]
    mode = os.environ.get("AB_MODE", mode)
    E = int(os.environ.get("AB_ENVS", E))
    if os.environ.get("AB_SHAPES"):
        shapes = [tuple(int(x) for x in s.split(",")) for s in os.environ["AB_SHAPES"].split(";")]
    rot = bench.Rotating(pkg, torch, dev, E, N, M, 30, 0)
    for g, th in shapes:
        try:
            rot.configure(g, th, mode)
            if os.environ.get("AB_VARIANT"):       # 1 = latency build (128 registers), 2 = occupancy build (64)
                for env in rot.envs:
                    pkg._lib.check(env.lib.chs_set_register_variant(env._h, int(os.environ["AB_VARIANT"])))
        except RuntimeError as exc:
            print(f"{tag} {name} G={g} T={th}: {exc}")
            continue
        rot.run(2 * rot.R)
        graph = rot.capture(K)
        graph.replay()
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            graph.replay()
            e.record()
            torch.cuda.synchronize()
            ts.append(1e3 * s.elapsed_time(e) / K)
        print(f"{tag} {name} E={E} {N}d/{M}c G={g} T={th} {mode}: median {np.median(ts):.2f} us/step  (min {min(ts):.2f}, max {max(ts):.2f}; R={rot.R})", flush=True)
    rot.close()
