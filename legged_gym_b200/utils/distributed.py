"""Env sharding across the GPUs of one box: the only exchange on this path.

Envs are independent; rank r owns global envs [r*N, (r+1)*N) with a full terrain
replica.  The only cross-env quantities are logging / curriculum reductions -- per-term episode
reward means over the envs reset, the mean terrain level, the reset count
(reference legged_robot.py:182-190) -- so one small all-reduce(sum) of the kernel's accumulated
statistics every `interval` steps is all the communication there is.  Means are formed after the
reduce, so they equal the single-process means up to fp32 summation order.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class ShardedStats:
    def __init__(self, env, interval: int = 24, group=None):
        self.env = env
        self.interval = max(int(interval), 1)
        self.group = group
        self.rows = env._num_sum_rows
        self.half = self.rows + 4
        self.global_extras = {}
        dev = env.episode_stats.device
        self.reduced = torch.zeros(self.half, dtype=torch.float, device=dev)
        # the all-reduce and the means run on a side stream so that env steps queued behind this call
        # do not wait for the other ranks; only the snapshot of the accumulators stays on the step stream
        self.side = torch.cuda.Stream(dev) if dev.type == "cuda" else None
        self._snap = torch.cuda.Event() if self.side is not None else None
        self._busy = torch.cuda.Event() if self.side is not None else None
        self._pending = False

    def after_step(self, step_index: int):
        """Call once per env.step(); every `interval` steps the accumulated sums are all-reduced."""
        if (step_index + 1) % self.interval != 0:
            return None
        return self.reduce_now()

    def reduce_now(self):
        acc = self.env.episode_stats[self.half:2 * self.half]
        if self.side is None:
            self.reduced.copy_(acc)
            acc[:self.rows + 1].zero_()
            if dist.is_available() and dist.is_initialized():
                dist.all_reduce(self.reduced, op=dist.ReduceOp.SUM, group=self.group)
            return self.publish()
        cur = torch.cuda.current_stream(self.reduced.device)
        if self._pending:
            cur.wait_event(self._busy)          # the previous reduce still owns self.reduced
        self.reduced.copy_(acc)
        acc[:self.rows + 1].zero_()
        self._snap.record(cur)
        with torch.cuda.stream(self.side):
            self.side.wait_event(self._snap)
            if dist.is_available() and dist.is_initialized():
                dist.all_reduce(self.reduced, op=dist.ReduceOp.SUM, group=self.group)
            out = self.publish()
            self._busy.record(self.side)
        self._pending = True
        return out

    def wait(self):
        """Make the current stream wait for the last reduce (before reading global_extras on it)."""
        if self._pending:
            torch.cuda.current_stream(self.reduced.device).wait_event(self._busy)

    def publish(self):
        """Global extras["episode"] values (device tensors; no host sync).  Two small kernels for the
        whole vector; the dict entries are views into it."""
        r = self.reduced
        means = r[:self.rows] / (r[self.rows].clamp(min=1.0) * self.env.max_episode_length_s)
        out = {'rew_' + name: means[row] for name, row in self.env._sum_row.items()}
        if self.env.cfg.terrain.curriculum:
            out["terrain_level"] = r[self.rows + 1] / r[self.rows + 2].clamp(min=1.0)
        out["num_resets"] = r[self.rows].clone()     # self.reduced is overwritten by the next snapshot
        self.global_extras = out
        return out
