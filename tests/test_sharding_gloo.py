"""N>1 path on CPU: world_size-2 gloo processes shard the env rows (no data-path
collective) and all-gather only the per-shard statistics."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank: int, world: int, port: int, n_envs: int, out_dir: str):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from conftest import PANDA
    from maniskill_b200.sharding import ShardStats, gather_stats, init_distributed, reduce_stats, shard_range
    from oracle import c_oracle as CO
    from oracle import oracle as O

    r, _, w = init_distributed(backend="gloo")
    assert (r, w) == (rank, world)
    fr = O.load_chain(PANDA[0], PANDA[1])
    rng = np.random.default_rng(2022)  # same global inputs on every rank
    q = rng.uniform(-1.5, -0.2, (n_envs, 7))
    a = rng.uniform(-0.1, 0.1, (n_envs, 6))
    lo, hi = shard_range(n_envs, rank, world)
    out = CO.delta_step(fr, q[lo:hi], a[lo:hi])  # the shard's work: no communication
    np.save(os.path.join(out_dir, f"shard{rank}.npy"), out)
    stats = ShardStats(solves=hi - lo, elapsed_ms=1.0 + rank, err_sum=float(np.abs(out).sum()), err_max=float(np.abs(out).max()),
                       iterations=hi - lo)
    allst = gather_stats(stats)
    agg = reduce_stats(allst)
    assert len(allst) == world and agg["solves"] == n_envs and agg["elapsed_ms"] == float(world)
    assert abs(agg["iterations_mean"] - 1.0) < 1e-12
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_env_sharding_gloo(tmp_path):
    n_envs, world = 1001, 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_envs, str(tmp_path)), nprocs=world, join=True)
    sys.path.insert(0, ROOT)
    from conftest import PANDA
    from oracle import c_oracle as CO
    from oracle import oracle as O

    fr = O.load_chain(PANDA[0], PANDA[1])
    rng = np.random.default_rng(2022)
    q = rng.uniform(-1.5, -0.2, (n_envs, 7))
    a = rng.uniform(-0.1, 0.1, (n_envs, 6))
    full = CO.delta_step(fr, q, a)
    parts = np.concatenate([np.load(tmp_path / f"shard{r}.npy") for r in range(world)])
    assert parts.shape == full.shape and np.array_equal(parts, full)
