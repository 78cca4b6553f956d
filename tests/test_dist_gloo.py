"""CPU, world_size 2 over gloo: the env-index sharding rule and the statistics reduction (the only collective)."""
import os
import subprocess
import sys
import textwrap

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_shard_range_partitions_every_env_once():
    from rl_cattle_herding_b200 import dist as chd
    for total, ws in ((65536, 8), (4096, 1), (10, 4), (7, 8), (65536, 3)):
        seen = []
        for r in range(ws):
            first, count = chd.shard_range(total, ws, r)
            seen.extend(range(first, first + count))
        assert seen == list(range(total))
        counts = [chd.shard_range(total, ws, r)[1] for r in range(ws)]
        assert max(counts) - min(counts) <= 1


WORKER = textwrap.dedent("""
    import os, sys, json
    import numpy as np
    import torch.distributed as dist
    sys.path.insert(0, %r)
    from rl_cattle_herding_b200 import dist as chd
    dist.init_process_group("gloo")
    rank, ws = chd.world()
    first, count = chd.shard_range(1000, ws, rank)
    stats = np.zeros(16); stats[0] = count; stats[1] = float(first); stats[10] = 7.0 * count
    red = chd.reduce_stats(stats)
    import torch
    obs = torch.full((count, 2), float(rank))
    allobs = chd.gather_observations(obs)
    # uneven shards (total %% world != 0): sizes are exchanged first, short shards padded for the collective
    f2, c2 = chd.shard_range(1001, ws, rank)
    uneven = chd.gather_observations(torch.full((c2, 3), float(rank)))
    if rank == 0:
        print(json.dumps({"red": red.tolist(), "ws": ws, "gathered": list(allobs.shape), "sum": chd.summarize(red)["episodes"],
                          "uneven": list(uneven.shape), "uneven_split": int((uneven[:, 0] == 0).sum())}))
    dist.barrier()
    dist.destroy_process_group()
""")


def test_stats_reduction_world_size_2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29731", str(script)],
                         capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    d = json.loads(line)
    assert d["ws"] == 2 and d["red"][0] == 1000 and d["red"][1] == 500.0 and d["red"][10] == 7000.0
    assert d["gathered"] == [1000, 2] and d["sum"] == 1000
    assert d["uneven"] == [1001, 3] and d["uneven_split"] == 501


def test_learners_refuse_uneven_shards():
    import pytest
    from rl_cattle_herding_b200 import dist as chd
    chd.require_even_shards(4096, 8)
    with pytest.raises(ValueError):
        chd.require_even_shards(1001, 2)


def test_bench_reference_arm_other_ranks_are_silent():
    """Under torchrun only rank 0 runs the CPU arm; the others exit 0 without output."""
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, env=env, timeout=120)
    assert out.returncode == 0 and out.stdout.strip() == ""


GRAD_WORKER = textwrap.dedent("""
    import os, sys, json
    import torch
    import torch.distributed as dist
    sys.path.insert(0, %r)
    from rl_cattle_herding_b200 import ppo
    dist.init_process_group("gloo")
    rank, ws = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)                                   # same weights on every rank
    pol = ppo.MlpActorCritic(6, 2, (8, 8), -1.0)
    torch.manual_seed(100 + rank)                          # different data per rank (its env shard)
    obs, act = torch.randn(32, 6), torch.randn(32, 2)
    mean, log_std = pol.dist(obs)
    loss = -pol.log_prob(mean, log_std, act).mean() + pol.value(obs).pow(2).mean()
    loss.backward()
    local = torch.cat([p.grad.reshape(-1) for p in pol.parameters()]).clone()
    ppo.average_gradients(list(pol.parameters()))
    avg = torch.cat([p.grad.reshape(-1) for p in pol.parameters()])
    both = [torch.zeros_like(local) for _ in range(ws)]
    dist.all_gather(both, local)
    expect = sum(both) / ws
    if rank == 0:
        print(json.dumps({"ws": ws, "err": float((avg - expect).abs().max()), "differs": float((both[0] - both[1]).abs().max())}))
    dist.barrier()
    dist.destroy_process_group()
""")


def test_learner_gradient_average_world_size_2(tmp_path):
    """Per-GPU learners (ppo.average_gradients): every rank ends up with the mean of the ranks' gradients."""
    script = tmp_path / "grad_worker.py"
    script.write_text(GRAD_WORKER % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29733", str(script)],
                         capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    d = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert d["ws"] == 2 and d["err"] < 1e-6 and d["differs"] > 1e-3
