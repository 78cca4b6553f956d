"""Training drivers on the GPU simulator: `python -m rl_cattle_herding_b200.train --mode ctde|dtde ...`.

CTDE mirrors `simulator/CTDECattleHerder.py` (one SB3-style MlpPolicy PPO over the whole observation), DTDE mirrors
`simulator/DTDECattleHerder.py` (shared-policy PPO over the per-drone streams of the MARL packing); see `ppo.py` for
what is and is not reproduced.  Under torchrun (one process per GPU) the envs are sharded by global env id and the
learners average gradients over NCCL.  Prints one JSON line per iteration and a final summary with the end-to-end
rate (rollout + update) next to the rollout-only and simulator-only rates.
"""
import argparse
import json
import os
import time

import torch


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="ctde", choices=["ctde", "dtde"])
    ap.add_argument("--envs", type=int, default=4096, help="total envs over all ranks")
    ap.add_argument("--drones", type=int, default=3)
    ap.add_argument("--cattle", type=int, default=20)
    ap.add_argument("--ctrl-freq", type=int, default=30)
    ap.add_argument("--level", type=int, default=None, help="curriculum start level (default: the reference class's)")
    ap.add_argument("--n-steps", type=int, default=16)
    ap.add_argument("--batch-size", type=int, default=0)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--no-graphs", action="store_true", help="eager PyTorch launches instead of CUDA graphs")
    args = ap.parse_args(argv)

    import torch.distributed as dist
    from . import spawn
    from .batched_env import CattleHerdingBatchedEnv
    from .dist import reduce_stats, shard_range
    from .ppo import DevicePPO

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from .dist import require_even_shards
    require_even_shards(args.envs, world)
    first, count = shard_range(args.envs, world, rank)
    marl = args.mode == "dtde"
    level = args.level if args.level is not None else (0 if marl else 7)       # MARLCattleAviary.py:62 / CattleAviary.py:62
    env = CattleHerdingBatchedEnv(count, num_drones=args.drones, num_cattle=args.cattle, ctrl_freq=args.ctrl_freq,
                                  fixed_num_drones=args.drones, start_level=level, auto_reset=True,
                                  spawn_table=spawn.generate(100, args.cattle, seed=0), device=local,
                                  env_id_offset=first, seed=args.seed, packing="marl" if marl else "ctde")
    algo = DevicePPO(env, n_steps=args.n_steps, batch_size=args.batch_size or None, seed=args.seed + rank,
                     cuda_graphs=False if args.no_graphs else None)
    if world > 1:      # same initial weights everywhere
        for p_ in algo.policy.parameters():
            dist.broadcast(p_.data, 0)

    def log(info):
        if rank == 0:
            print(json.dumps({k: (round(v, 6) if isinstance(v, float) else v) for k, v in info.items()}), flush=True)

    hist = algo.learn(args.iters, log=log)

    # simulator alone on the same shapes, policy out of the loop (random actions regenerated on device)
    acts = torch.rand((count, args.drones, env.action_width), device=env.device) * 2 - 1
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n_sim = 200
    for _ in range(n_sim):
        env.step(acts)
    torch.cuda.synchronize()
    sim_rate = n_sim * count / (time.perf_counter() - t0)
    stats = reduce_stats(env.stats())
    if world > 1:
        t = torch.tensor([sim_rate, hist[-1]["env_steps_per_s"], hist[-1]["rollout_env_steps_per_s"]], device=env.device,
                         dtype=torch.float64)
        dist.all_reduce(t)
        sim_rate, e2e, roll = t.tolist()
    else:
        e2e, roll = hist[-1]["env_steps_per_s"], hist[-1]["rollout_env_steps_per_s"]
    if rank == 0:
        print(json.dumps({"summary": args.mode, "n_gpus": world, "total_envs": args.envs, "n_steps": args.n_steps,
                          "samples_per_iteration": algo.n_steps * algo.B * world, "minibatch": algo.batch_size, "cuda_graphs": algo.cuda_graphs,
                          "train_env_steps_per_s": e2e, "rollout_env_steps_per_s": roll, "sim_only_env_steps_per_s": sim_rate,
                          "agent_steps_per_s": e2e * args.drones, "episodes_seen": float(stats[0]),
                          "hyperparameters": {k: v for k, v in algo.hp.items()}}), flush=True)
    env.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
