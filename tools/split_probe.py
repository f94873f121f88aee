#!/usr/bin/env python
"""How long do the rules alone and the encode alone take?  (profiling aid)
    python tools/split_probe.py [workload]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import bench  # noqa: E402
from microrts_py_b200 import _capi, microrts_ai  # noqa: E402
from microrts_py_b200.envs import MicroRTSGridModeVecEnv  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "diag_wrush16"
wl = bench.workload_setup(name, 0)
env = MicroRTSGridModeVecEnv(wl["nsp"], wl["nbot"], partial_obs=wl["partial"], max_steps=wl["max_steps"], ai2s=[getattr(microrts_ai, b) for b in wl["bot_names"]],
                             map_paths=[wl["map"]], device=0, return_torch=True, seed=1000)
L, N, H, W_, P = env._lib, env.num_envs, env.height, env.width, sum(env.num_planes)
dev = env.device
obs = torch.empty((N, H, W_, P), dtype=torch.int32, device=dev)
mask = torch.empty((N, H * W_, 78), dtype=torch.int32, device=dev)
rew = torch.empty((N, 6), dtype=torch.float64, device=dev)
done = torch.empty((N,), dtype=torch.uint8, device=dev)
act = torch.zeros((N, H * W_, 7), dtype=torch.int32, device=dev)
sp = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
io = _capi.StepIOC()
io.actions, io.obs, io.mask, io.reward, io.done, io.next_actions, io.seed = act.data_ptr(), obs.data_ptr(), mask.data_ptr(), rew.data_ptr(), done.data_ptr(), act.data_ptr(), 7
_capi.check(L.mrts_reset(env._h, obs.data_ptr(), mask.data_ptr(), sp))
_capi.check(L.mrts_sample_actions(env._h, act.data_ptr(), C.c_uint64(7), sp))
for _ in range(300):
    _capi.check(L.mrts_step_ex(env._h, C.byref(io), sp))


def timed(fn, n=30):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


fused = timed(lambda: _capi.check(L.mrts_step_ex(env._h, C.byref(io), sp)))
io2 = _capi.StepIOC()
io2.actions, io2.obs, io2.mask, io2.reward, io2.done, io2.next_actions, io2.seed = act.data_ptr(), 0, 0, rew.data_ptr(), done.data_ptr(), act.data_ptr(), 7
rules = timed(lambda: _capi.check(L.mrts_step_ex(env._h, C.byref(io2), sp)))
enc = timed(lambda: _capi.check(L.mrts_reset(env._h, obs.data_ptr(), mask.data_ptr(), sp)))  # template load + obs + mask of every game
print(f"{name}: fused step {fused:.3f} ms | step without obs/mask outputs {rules:.3f} ms | reset (no rules, obs + mask written) {enc:.3f} ms")
env.close()
