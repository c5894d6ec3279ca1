"""CPU tier: the N>1 path -- env sharding + the one all-reduce of episode statistics -- on 2 gloo ranks.

The kernels fill env.episode_stats (second half = sums accumulated since the last reduce); here the ranks
write known values there, so the test covers the host-side protocol: what is reduced, how the means are
formed, that accumulators are cleared, and that every rank ends with the same global extras."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import helpers
        from legged_gym_b200.utils.distributed import ShardedStats

        def shard(cfg):
            cfg.env.shard_rank, cfg.env.shard_world = rank, world
        env = helpers.make_env("go1_rough_full", 32, extra=shard)
        R = env._num_sum_rows
        st = ShardedStats(env, interval=3)
        half = R + 4
        # what the step kernel would have accumulated on this rank: sums over reset envs, count, levels, N
        sums = torch.arange(R, dtype=torch.float) * (rank + 1)
        n_reset = float(3 + rank)
        level_sum = float(env.terrain_levels.sum())
        env.episode_stats[half:half + R] = sums
        env.episode_stats[half + R] = n_reset
        env.episode_stats[half + R + 1] = level_sum
        env.episode_stats[half + R + 2] = 32
        assert st.after_step(0) is None and st.after_step(1) is None      # only every `interval` steps
        ex = st.after_step(2)
        assert float(env.episode_stats[half:half + R + 1].abs().sum()) == 0.0   # accumulators cleared
        res = {k: float(v) for k, v in ex.items()}
        # the command curriculum's gate: mean over the reset envs of ALL ranks (rank 0: 2 envs at 10, rank 1: 4 at 40)
        vals = torch.full((2 + 2 * rank,), 10.0 + 30.0 * rank)
        res["_gate_mean"] = float(env._mean_over_all_shards(vals))
        assert env._mean_over_all_shards(torch.zeros(0)) is None and env._shard_world() == world
        res["_types"] = env.terrain_types.tolist()
        res["_levels"] = level_sum
        out[rank] = res
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_stats_allreduce_world2():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    a, b = out[0], out[1]
    keys = [k for k in a if not k.startswith("_")]
    for k in keys:
        assert a[k] == b[k], f"ranks disagree on {k}"
    # global mean = (sum over ranks) / (resets over ranks) / max_episode_length_s(20)
    assert a["num_resets"] == 7.0
    assert a["_gate_mean"] == b["_gate_mean"] == (2 * 10.0 + 4 * 40.0) / 6
    assert abs(a["rew_ang_vel_xy"] - (1 * 1 + 1 * 2) / 7.0 / 20.0) < 1e-6
    assert abs(a["terrain_level"] - (a["_levels"] + b["_levels"]) / 64.0) < 1e-6
    # terrain types partition the global index range: rank 0 owns the low types, rank 1 the high ones
    assert max(a["_types"]) <= min(b["_types"])
