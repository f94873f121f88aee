#!/usr/bin/env python
"""Per-game cycle histogram of one launch (profiling aid, mrts_debug_timing): shows whether a launch
is bounded by the sum of the games' work or by its slowest games.
    python tools/game_timing.py [workload] [prewarm]"""
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
prewarm = int(sys.argv[2]) if len(sys.argv) > 2 else 300
wl = bench.workload_setup(name, 0)
env = MicroRTSGridModeVecEnv(wl["nsp"], wl["nbot"], partial_obs=wl["partial"], max_steps=wl["max_steps"], ai2s=[getattr(microrts_ai, b) for b in wl["bot_names"]],
                             map_paths=[wl["map"]], device=0, return_torch=True, seed=1000)
L = env._lib
G = L.mrts_num_games(env._h)
obs = env.reset()
for _ in range(prewarm):
    m = env.get_action_mask()
    obs, r, d, _ = env.step(env.sample_actions())
tim = torch.zeros((G, 16), dtype=torch.int64, device=env.device)
_capi.check(L.mrts_debug_timing(env._h, C.c_void_p(tim.data_ptr())))
hist = []
for k in range(8):
    m = env.get_action_mask()
    obs, r, d, _ = env.step(env.sample_actions())
    torch.cuda.synchronize()
    hist.append(tim.cpu().numpy().copy())
_capi.check(L.mrts_debug_timing(env._h, None))
t = np.stack(hist).astype(np.float64)  # (8, G, 2)
names = ["rules", "total", "decode", "bot-puts", "bot-translate", "issue-agent", "issue-bot", "cycle", "#searches", "#expansions", "", "", "astar", "astar-guard", "#guards", "staging"]
for lbl, x in [(names[k], t[:, :, k]) for k in list(range(10)) + [12, 13, 14, 15]]:
    q = np.percentile(x, [50, 90, 99, 99.9, 100])
    print(f"{name} {lbl:14s} /game: mean {x.mean():9.0f}  p50 {q[0]:9.0f}  p90 {q[1]:9.0f}  p99 {q[2]:9.0f}  p99.9 {q[3]:9.0f}  max {q[4]:9.0f}")
# timeline of the last launch: resident games per SM over time, and how the launch ends
last = hist[-1].astype(np.float64)
start = last[:, 10] - last[:, 10].min()
dur = last[:, 1] / 1.965  # cycles -> ns at 1965 MHz
end = start + dur
T = end.max()
print(f"launch span {T/1e3:.1f} us; sum of game time {dur.sum()/1e6:.2f} ms -> mean concurrency {dur.sum()/T:.0f} games = {dur.sum()/T/148:.1f} per SM")
for frac in (0.1, 0.25, 0.5, 0.75, 0.9, 0.95, 0.99):
    t = frac * T
    print(f"  at {frac:4.0%} of the span: {int(((start <= t) & (end > t)).sum()):6d} games resident, {int((start <= t).sum()):6d} started")
slow = np.argsort(-last[:, 1])[:12]
print("slowest games of the last launch: total | rules | translate | astar | guard | #searches | #expansions | #guards | staging")
for g in slow:
    print("   ", " | ".join(str(int(last[g, k])) for k in (1, 0, 4, 12, 13, 8, 9, 14, 15)))
order = np.argsort(start)
print("  cost class of the first 2000 started (mean cycles)", last[order[:2000], 1].mean(), " of the last 2000", last[order[-2000:], 1].mean())
state = [env.export_state(g)[1] for g in range(0, G, max(1, G // 256))]
print("units/game mean", np.mean([len(s) for s in state]), "max", max(len(s) for s in state))
env.close()
