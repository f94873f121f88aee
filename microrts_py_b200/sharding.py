"""Env sharding across the GPUs of one box: one process per GPU, one independent slab of
environments per process, no collective in the step path.

The reference has no multi-process code; its vec client is a sequential loop over independent
per-env clients (J/tests/JNIGridnetVecClient.java:241-287), which is what makes the split
trivial.  The global env order is the reference's (JNIGridnetVecClient.java:117-124): the
``num_selfplay_envs`` player views first (game g = envs 2g, 2g+1), then the bot envs.  A shard
keeps that order locally, so rank r's local env i is global env ``Shard.global_index[i]`` and
concatenating the ranks' outputs through ``Shard.global_index`` restores the single-process
result exactly.  The two views of a self-play game always land on the same rank.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence

import numpy as np


def _split(n: int, world: int, rank: int):
    """Contiguous block [lo, hi) of n items for `rank`; the first n % world ranks get one more."""
    q, r = divmod(n, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


@dataclass
class Shard:
    rank: int
    world_size: int
    num_selfplay_envs: int      # local player views (even)
    num_bot_envs: int           # local bot envs
    ai2s: list                  # local slice of the global ai2s
    map_paths: list             # one entry, or one per LOCAL env in local order
    global_index: np.ndarray    # (local N,) global env index of each local env

    @property
    def num_envs(self) -> int:
        return self.num_selfplay_envs + self.num_bot_envs

    def ctor_kwargs(self) -> dict:
        return dict(num_selfplay_envs=self.num_selfplay_envs, num_bot_envs=self.num_bot_envs, ai2s=self.ai2s, map_paths=self.map_paths)


def shard_envs(num_selfplay_envs: int, num_bot_envs: int, world_size: int, rank: int, ai2s: Sequence = (), map_paths: Sequence[str] = ()) -> Shard:
    """The slab of rank `rank`: a contiguous block of self-play GAMES and a contiguous block of bot envs."""
    if num_selfplay_envs % 2:
        raise ValueError("num_selfplay_envs must be even")
    if not 0 <= rank < world_size:
        raise ValueError("rank out of range")
    if len(ai2s) != num_bot_envs:
        raise ValueError("for each bot environment, a microrts ai should be provided")
    n = num_selfplay_envs + num_bot_envs
    g_lo, g_hi = _split(num_selfplay_envs // 2, world_size, rank)
    b_lo, b_hi = _split(num_bot_envs, world_size, rank)
    idx = np.concatenate([np.arange(2 * g_lo, 2 * g_hi), num_selfplay_envs + np.arange(b_lo, b_hi)]).astype(np.int64)
    paths: List[str] = list(map_paths)
    if len(paths) > 1:
        if len(paths) != n:
            raise ValueError("if multiple maps are provided, they should be provided for each environment")
        paths = [paths[i] for i in idx]
    return Shard(rank, world_size, 2 * (g_hi - g_lo), b_hi - b_lo, list(ai2s[b_lo:b_hi]), paths, idx)


def shard_sizes(num_selfplay_envs: int, num_bot_envs: int, world_size: int) -> List[int]:
    """Local env count of every rank under the split `shard_envs` uses."""
    out = []
    for k in range(world_size):
        g_lo, g_hi = _split(num_selfplay_envs // 2, world_size, k)
        b_lo, b_hi = _split(num_bot_envs, world_size, k)
        out.append(2 * (g_hi - g_lo) + (b_hi - b_lo))
    return out


def gather_rewards_dones(reward, done, shard: Shard, num_selfplay_envs: int, num_bot_envs: int):
    """OFF the step path: all-gather of the (reward, done) scalars (9 B/env) for a central logger.
    Works with any initialised torch.distributed backend (NCCL on the GPUs, gloo in the CPU tests);
    returns (reward (total,), done (total,)) in the global env order on every rank."""
    import torch
    import torch.distributed as dist

    total = num_selfplay_envs + num_bot_envs
    sizes = shard_sizes(num_selfplay_envs, num_bot_envs, shard.world_size)
    r = torch.as_tensor(reward, dtype=torch.float64)
    dev = r.device
    pack = torch.zeros((max(sizes), 3), dtype=torch.float64, device=dev)
    pack[: shard.num_envs, 0] = r
    pack[: shard.num_envs, 1] = torch.as_tensor(done).to(device=dev, dtype=torch.float64)
    pack[: shard.num_envs, 2] = torch.as_tensor(shard.global_index, dtype=torch.float64, device=dev)
    out = [torch.empty_like(pack) for _ in range(shard.world_size)]
    dist.all_gather(out, pack)
    R = torch.zeros(total, dtype=torch.float64, device=dev)
    D = torch.zeros(total, dtype=torch.bool, device=dev)
    for k, o in enumerate(out):
        o = o[: sizes[k]]
        gi = o[:, 2].long()
        R[gi] = o[:, 0]
        D[gi] = o[:, 1] != 0
    return R, D
