"""Multi-GPU layout: one process per GPU, independent env slabs, no collective in the
step path.  The N>1 host logic is exercised here with world_size-2 `gloo` processes on CPU, each
stepping ITS slab on the host build of the device headers (tests/hostsim) and comparing against
the single-process run of the whole env list; the off-path (reward, done) all-gather is checked
too.  On the B200 box the same split runs one MicroRTSGridModeVecEnv per rank (bench.py --gpus N)."""
import os
import socket
import sys

import numpy as np
import pytest

from microrts_py_b200 import microrts_ai
from microrts_py_b200.sharding import shard_envs, shard_sizes
from tests.common import MAP_8, MAP_16

P, W = microrts_ai.passiveAI, microrts_ai.workerRushAI


def test_shards_partition_the_env_list():
    for nsp, nbot, world in [(0, 24, 8), (6, 5, 2), (2, 0, 4), (64, 3, 8), (10, 10, 3)]:
        ais = [P if i % 2 else W for i in range(nbot)]
        paths = [f"m{i}" for i in range(nsp + nbot)]
        seen = []
        for r in range(world):
            s = shard_envs(nsp, nbot, world, r, ais, paths)
            assert s.num_selfplay_envs % 2 == 0 and s.num_envs == len(s.global_index)
            assert s.map_paths == [paths[i] for i in s.global_index]
            assert s.ai2s == [ais[i - nsp] for i in s.global_index if i >= nsp]
            sp = s.global_index[: s.num_selfplay_envs]
            assert all(a % 2 == 0 and b == a + 1 for a, b in zip(sp[::2], sp[1::2]))  # pairs stay together
            seen.extend(s.global_index.tolist())
        assert sorted(seen) == list(range(nsp + nbot))
        assert shard_sizes(nsp, nbot, world) == [shard_envs(nsp, nbot, world, r, ais, paths).num_envs for r in range(world)]
    with pytest.raises(ValueError):
        shard_envs(3, 0, 2, 0)
    with pytest.raises(ValueError):
        shard_envs(2, 2, 2, 0, [P])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nsp, nbot, steps, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist

    from microrts_py_b200.sharding import gather_rewards_dones
    from tests.backends import make_env
    from tests.common import sample_actions

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ais = [P if i % 2 else W for i in range(nbot)]
        paths = [MAP_16 if (i // 2) % 2 == 0 else MAP_16 for i in range(nsp + nbot)]
        sh = shard_envs(nsp, nbot, world, rank, ais, paths)
        env = make_env("hostsim", max_steps=60, **sh.ctor_kwargs())
        full = make_env("hostsim", nsp, nbot, max_steps=60, ai2s=ais, map_paths=paths) if rank == 0 else None
        rng = np.random.default_rng(11)  # same stream on every rank: the GLOBAL action grid
        obs = env.reset()
        if full is not None:
            np.testing.assert_array_equal(full.reset()[sh.global_index], obs)
        ok = True
        for t in range(steps):
            m = env.get_action_mask()
            # global mask assembled only to drive the shared sampler (test plumbing, not the step path)
            parts = [None] * world
            dist.all_gather_object(parts, (sh.global_index, m))
            gm = np.zeros((nsp + nbot,) + m.shape[1:], m.dtype)
            for gi, mm in parts:
                gm[gi] = mm
            a = sample_actions(gm, rng, 0.03)
            o, r, d, infos = env.step(a[sh.global_index])
            R, D = gather_rewards_dones(r, d, sh, nsp, nbot)
            if full is not None:
                np.testing.assert_array_equal(full.get_action_mask(), gm)
                fo, fr, fd, _ = full.step(a)
                np.testing.assert_array_equal(fo[sh.global_index], o)
                np.testing.assert_array_equal(R.numpy(), fr)
                np.testing.assert_array_equal(D.numpy(), fd)
        dist.barrier()
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_slabs_equal_the_single_process_run():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    world, nsp, nbot, steps = 2, 6, 5, 90
    procs = [ctx.Process(target=_worker, args=(r, world, port, nsp, nbot, steps, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    got = sorted(q.get(timeout=5) for _ in range(world))
    assert got == [(0, True), (1, True)]
