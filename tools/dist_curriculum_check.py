"""2-GPU check of the command curriculum's global gate (run under torchrun, one rank per GPU):
rank 0's reset envs track well (episode sum 30), rank 1's badly (3); the mean over both ranks (16.5) is above
the gate (16), so BOTH ranks must widen lin_vel_x -- a per-rank decision would widen on rank 0 only.
usage: gpurun --gpus 2 -- 'python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
       --master-port 29533 tools/dist_curriculum_check.py'"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import helpers  # noqa: E402
from legged_gym_b200.utils import synthetic  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def cfg_edit(cfg):
    cfg.commands.curriculum = True
    cfg.commands.max_curriculum = 2.0
    cfg.env.shard_rank, cfg.env.shard_world = rank, world


env = helpers.make_env("go1_rough_full", 256, device=f"cuda:{local}", seed=1 + rank, extra=cfg_edit)
helpers.randomise_env_state(env, rank)
env.common_step_counter = int(env.max_episode_length) - 1
env.episode_sums["tracking_lin_vel"][:] = 30.0 if rank == 0 else 3.0
env.episode_length_buf[:] = 5
env.episode_length_buf[:16] = int(env.max_episode_length)          # 16 time-outs per rank
fr = synthetic.make_frames(env, 1, seed=7 + rank)[0]
fr["contact_forces"].zero_()                                       # no contact terminations: equal reset counts
synthetic.load_frame(env, fr)
env.step(fr["actions"].clone())
torch.cuda.synchronize()
got = [float(v) for v in env.command_ranges["lin_vel_x"]]
n = int(env.step_counts[0])
print(f"rank {rank}: {n} resets, lin_vel_x range after the gated step {got}", flush=True)
assert n == 16 and got == [-1.5, 1.5], (n, got)
dist.barrier()
if rank == 0:
    print("global command-curriculum gate OK on", world, "ranks")
dist.destroy_process_group()
