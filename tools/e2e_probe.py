#!/usr/bin/env python
"""Where an end-to-end step (bench.py e2e loop) spends its time: policy stand-in, its action grid
device->host, env.step (mrts_step_host_slot).  Usage: python tools/e2e_probe.py [workload] [envs] [prewarm steps]"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from microrts_py_b200 import microrts_ai  # noqa: E402
from microrts_py_b200.envs import MicroRTSGridModeVecEnv  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg2_rollout16"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
prewarm = int(sys.argv[3]) if len(sys.argv) > 3 else 5
wl = bench.workload_setup(name, None)
nsp = int(n * (wl["nsp"] / wl["n"])) // 2 * 2
nbot = n - nsp
ai2s = [getattr(microrts_ai, wl["bot_names"][i * wl["nbot"] // nbot]) for i in range(nbot)] if nbot else []
env = MicroRTSGridModeVecEnv(nsp, nbot, partial_obs=wl["partial"], max_steps=wl["max_steps"], ai2s=ai2s, map_paths=[wl["map"]], host_transport=True, seed=3)
a_host = torch.zeros((n, env.height * env.width, 7), dtype=torch.int32).pin_memory()
env.reset()
T = {"sample": 0.0, "actions_d2h": 0.0, "step": 0.0}
K = 40
lib_s = [0.0]
_orig = env._lib.mrts_step_host_slot


def _timed(*args):
    t = time.perf_counter()
    r = _orig(*args)
    lib_s[0] += time.perf_counter() - t
    return r


env._lib.mrts_step_host_slot = _timed
for it in range(K + prewarm):
    if it == prewarm:
        lib_s[0] = 0.0
        T = dict.fromkeys(T, 0.0)
        ph = (C.c_double * 8)()
        env._lib.mrts_host_timing(env._h, ph, 1)
    t0 = time.perf_counter()
    a = env.sample_actions()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    a_host.copy_(a, non_blocking=True)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    env.step(a_host.numpy())
    env.get_action_mask()
    t3 = time.perf_counter()
    T["sample"] += t1 - t0
    T["actions_d2h"] += t2 - t1
    T["step"] += t3 - t2
tot = sum(T.values())
print(f"{name} {n} envs: " + " | ".join(f"{k} {1e3 * v / K:.2f} ms" for k, v in T.items()) + f" | total {1e3 * tot / K:.2f} ms = {n * K / tot / 1e6:.2f} M env-steps/s")
ph = (C.c_double * 8)()
env._lib.mrts_host_timing(env._h, ph, 0)
calls = max(ph[7], 1)
names = ["pack actions", "h2d + launch", "wait device", "patch + copy-out"]
print("  inside mrts_step_host_slot (%d threads): " % env._lib.mrts_host_threads(env._h) + " | ".join(f"{k} {1e3 * ph[i] / calls:.3f} ms" for i, k in enumerate(names))
      + f" | {ph[5] / calls / n:.2f} changed cells/env, {ph[6] / calls / 1e6:.2f} MB d2h, {ph[4] / calls / 1e6:.2f} MB h2d" + f" | inside the C call {1e3 * lib_s[0] / K:.3f} ms" + f" | env.step total {1e3 * T['step'] / K:.3f} ms = {n * K / T['step'] / 1e6:.2f} M env-steps/s")
try:
    roll = dict(l.split(":") for l in open("/proc/self/smaps_rollup").read().splitlines()[1:] if ":" in l)
    print("  process memory: AnonHugePages", roll.get("AnonHugePages", "?").strip(), "of Anonymous", roll.get("Anonymous", "?").strip(),
          "| THP", open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip(), "defrag", open("/sys/kernel/mm/transparent_hugepage/defrag").read().strip())
except Exception as ex:  # noqa: BLE001
    print("  (no smaps_rollup:", ex, ")")
