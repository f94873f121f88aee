#!/usr/bin/env python
"""Throughput of the vectorised MicroRTS environment step (BASELINE.json metric: env-steps/s
including observation + mask + reward, and achieved HBM GB/s against the measured peak).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--envs E] [--dtype int32|uint8]
  python bench.py --impl reference ...      # the CPU restatement of the reference path (oracle/) on host cores

One "step" = one pass of the hot path over one slab of environments: mrts_step_ex =
JNIGridnetVecClient.gameStep fused with getMasks, writing obs (N,H,W,P), mask (N,HW,78), raw
rewards (N,6) and dones (N,) to HBM.  One env-step = one player view advanced one frame.
`value` times that with actions already resident in HBM (drawn on the device by the fused
masked-uniform sampler); `e2e` times the reference-facing class with HOST buffers: actions
host->device, observation / mask / rewards / dones device->host every step.
Multi-GPU: one process per GPU (torchrun), independent slabs, no collective in the step path
(weak scaling); NCCL is used only for the start/stop barrier and the max-over-ranks time.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from tests.common import MAP_8, MAP_16, MAP_32, MAP_64  # noqa: E402  (map names only)

# name: (map, planes, envs per GPU, fraction self-play, bots, partial_obs, max_steps, description)
WORKLOADS = {
    "cfg2_rollout16": (MAP_16, 29, 24576, 0.0, ("passiveAI", "workerRushAI"), False, 2000,
                       "ppo_gridnet rollout shape x1024: 16x16 basesWorkers16x16A, bot envs, half vs passiveAI half vs workerRushAI"),
    "diag_passive16": (MAP_16, 29, 24576, 0.0, ("passiveAI",), False, 2000, "diagnostic: 16x16 bot envs, all vs passiveAI"),
    "diag_wrush16": (MAP_16, 29, 24576, 0.0, ("workerRushAI",), False, 2000, "diagnostic: 16x16 bot envs, all vs workerRushAI"),
    "diag_lrush16": (MAP_16, 29, 24576, 0.0, ("lightRushAI",), False, 2000, "diagnostic: 16x16 bot envs, all vs lightRushAI"),
    "diag_mix3_16": (MAP_16, 29, 24576, 0.0, ("passiveAI", "workerRushAI", "lightRushAI"), False, 2000, "diagnostic: 16x16 bot envs, thirds vs passiveAI / workerRushAI / lightRushAI"),
    "cfg1_selfplay16": (MAP_16, 29, 32768, 1.0, (), False, 2000, "16x16 basesWorkers16x16A self-play"),
    "cfg3_selfplay8": (MAP_8, 29, 65536, 1.0, (), False, 2000, "8x8 basesWorkers8x8 self-play, 65,536 envs per GPU"),
    "cfg4_partial32": (MAP_32, 31, 16384, 1.0, (), True, 2000, "32x32 BWDistantResources32x32 self-play, partial_obs, 16,384 envs per GPU"),
    "cfg5_bloodbath64": (MAP_64, 29, 2048, 1.0, (), False, 2000, "64x64 (4)BloodBath.scmB self-play"),
}
DEFAULT_WORKLOAD = "cfg2_rollout16"
# dram__bytes_read.sum + dram__bytes_write.sum of the pass kernel, bytes per launch, from the ncu --set full captures
# under profiles/ (same command line, same slab size); None where no capture exists
NCU_TRAFFIC = {
    ("cfg2_rollout16", "int32"): 26_905_856 + 2_825_127_000,   # profiles/ncu_r01_cfg2_details.csv (read + write)
    ("cfg1_selfplay16", "int32"): 23_728_640 + 3_705_619_000,  # profiles/ncu_r01_cfg1_details.csv
}


_RESULT_OUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout; libraries (NCCL's version banner, torchrun notices) also
    write to fd 1.  Keep a private handle to the real stdout for the result line and point fd 1 at stderr."""
    global _RESULT_OUT
    if _RESULT_OUT is None:
        sys.stdout.flush()
        _RESULT_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _RESULT_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def algorithmic_bytes(hw, planes, elem):
    """B = obs + mask + actions_in + 6*8 + 1 per env-step, at the stored dtypes."""
    return hw * planes * elem + hw * 78 * elem + hw * 7 * 4 + 49


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:  # noqa: BLE001
                self.proc.kill()
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def workload_setup(name, envs):
    mp, planes, n_default, frac_sp, bots, partial, max_steps, desc = WORKLOADS[name]
    n = envs or n_default
    nsp = int(n * frac_sp) // 2 * 2
    nbot = n - nsp
    bot_names = [bots[i * len(bots) // nbot] for i in range(nbot)] if nbot else []
    return dict(map=mp, planes=planes, n=n, nsp=nsp, nbot=nbot, bot_names=bot_names, partial=partial, max_steps=max_steps, desc=desc)


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (oracle/, a C++ restatement of the reference's Java path -- the JVM/JPype
# reference itself cannot run on this image) on the host cores.  The only place besides tests/ and
# smoke() that executes oracle/ code, and only as the thing compared against, never shipped.
# ------------------------------------------------------------------------------------------------
def cpu_run(wl, n_envs, steps, warmup, budget_s, threads):
    from oracle.oracle_py import OracleVecEnv
    from tests.common import sample_actions, templates_for

    nsp = int(n_envs * (wl["nsp"] / wl["n"])) // 2 * 2
    nbot = n_envs - nsp
    ids = [{"passiveAI": 0, "workerRushAI": 1, "lightRushAI": 2}[wl["bot_names"][i * wl["nbot"] // nbot]] for i in range(nbot)] if nbot else []
    env = OracleVecEnv(nsp, nbot, wl["max_steps"], wl["partial"], templates_for([wl["map"]] * n_envs), ids, threads=threads)
    rng = np.random.default_rng(0)
    obs = env.reset()
    out = (obs, np.zeros((n_envs, 6)), np.zeros(n_envs, np.uint8))
    spent, done_steps, per_step = 0.0, 0, []
    wall0 = time.perf_counter()
    for t in range(warmup + steps):
        t0 = time.perf_counter()
        m, _ = env.get_masks()
        t1 = time.perf_counter()
        a = sample_actions(m, rng)  # the policy stand-in: NOT timed
        t2 = time.perf_counter()
        env.step(a, out)
        t3 = time.perf_counter()
        if t >= warmup:
            dt = (t1 - t0) + (t3 - t2)
            spent += dt
            per_step.append(dt)
            done_steps += 1
            # the budget bounds the leg's WALL time: the untimed NumPy action sampling costs ~10x the timed calls
            if budget_s and time.perf_counter() - wall0 > budget_s:
                break
    return n_envs * done_steps / spent, done_steps, spent


def cpu_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        return os.cpu_count() or 1


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = workload_setup(args.workload, args.envs)
    cores = cpu_cores()
    n_cpu = args.cpu_envs
    v, k, spent = cpu_run(wl, n_cpu, args.steps, args.warmup, 150.0, cores)
    hw = {MAP_8: 64, MAP_16: 256, MAP_32: 1024, MAP_64: 4096}[wl["map"]]
    sample = f"{n_cpu} envs of the workload per step, {k} timed steps ({spent:.1f} s), get_masks + step timed, action sampling excluded"
    line = {
        "impl": "reference", "metric": "env-steps/sec incl. obs+mask+reward", "value": v, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": k, "warmup": args.warmup, "ms_per_step": 1e3 * spent / max(k, 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic (real reference maps, seeded masked-uniform actions)",
        "config": {"workload": args.workload, "description": wl["desc"], "map": wl["map"], "envs_per_step": n_cpu, "cells": hw},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample,
                         "note": "C++ restatement of the reference's Java path (oracle/); the JVM/JPype reference is unavailable on this box"},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def gpu_arm(args):
    import ctypes as C

    import torch

    from microrts_py_b200 import _capi, microrts_ai
    from microrts_py_b200.envs import MicroRTSGridModeVecEnv

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    wl = workload_setup(args.workload, args.envs)
    N, nsp, nbot = wl["n"], wl["nsp"], wl["nbot"]
    np_dtype = np.int32 if args.dtype == "int32" else np.uint8
    elem = np.dtype(np_dtype).itemsize
    ai2s = [getattr(microrts_ai, b) for b in wl["bot_names"]]
    env = MicroRTSGridModeVecEnv(nsp, nbot, partial_obs=wl["partial"], max_steps=wl["max_steps"], ai2s=ai2s, map_paths=[wl["map"]],
                                 device=local, dtype=np_dtype, return_torch=True, seed=1000 + rank)
    H, Wd, Pn = env.height, env.width, sum(env.num_planes)
    HW = H * Wd
    L = env._lib
    dev = env.device
    t_dtype = torch.int32 if elem == 4 else torch.uint8
    obs = torch.empty((N, H, Wd, Pn), dtype=t_dtype, device=dev)
    mask = torch.empty((N, HW, 78), dtype=t_dtype, device=dev)
    rew = torch.empty((N, 6), dtype=torch.float64, device=dev)
    done = torch.empty((N,), dtype=torch.uint8, device=dev)
    actions = torch.zeros((N, HW, 7), dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)
    sp = C.c_void_p(stream.cuda_stream)

    io = _capi.StepIOC()
    io.actions, io.obs, io.mask, io.reward, io.done = actions.data_ptr(), obs.data_ptr(), mask.data_ptr(), rew.data_ptr(), done.data_ptr()
    io.next_actions, io.seed = actions.data_ptr(), 1000 + rank

    def step():
        _capi.check(L.mrts_step_ex(env._h, C.byref(io), sp))

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    _capi.check(L.mrts_reset(env._h, obs.data_ptr(), mask.data_ptr(), sp))
    _capi.check(L.mrts_sample_actions(env._h, actions.data_ptr(), C.c_uint64(1000 + rank), sp))
    # spread the games over episode phases so that the timed steps see a mix of early / mid / late states
    for _ in range(args.prewarm):
        step()
    for _ in range(args.warmup):
        step()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    dones = torch.zeros((), dtype=torch.int64, device=dev)
    ev[0].record(stream)
    for k in range(args.steps):
        step()
        ev[k + 1].record(stream)
    barrier()
    clk = clocks.stop() if rank == 0 else None
    total_ms = ev[0].elapsed_time(ev[-1])
    per_launch = [ev[k].elapsed_time(ev[k + 1]) for k in range(args.steps)]
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    units_mean = float(mask[:, :, 0].sum().item()) / N  # units that may act per view, end state

    # ---------------- end to end: the reference-facing class with HOST buffers
    e2e = e2e_run(args, env, wl, rank, dist, dev) if args.e2e_envs > 0 else None

    if rank == 0:
        B = algorithmic_bytes(HW, Pn, elem)
        peak, peak_src = measured_peak()
        kern_ms = float(np.mean(per_launch))
        achieved = B * N / (kern_ms * 1e-3) / 1e9
        value = N * world * args.steps / (total_ms_max * 1e-3)
        cores = cpu_cores()
        cpu_v, cpu_k, cpu_spent = cpu_run(wl, args.cpu_envs, 10 ** 9, 1, args.cpu_budget, cores) if (world == 1 and args.cpu_budget > 0) else (None, 0, 0.0)
        line = {
            "metric": "env-steps/sec incl. obs+mask+reward", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic (real reference maps; actions drawn on the device by the fused masked-uniform sampler)",
            "config": {"workload": args.workload, "description": wl["desc"], "map": wl["map"], "envs_per_gpu": N, "selfplay_views": nsp, "bot_envs": nbot,
                       "partial_obs": wl["partial"], "max_steps": wl["max_steps"], "prewarm_steps": args.prewarm, "l2": "outputs per step (%.2f GB) exceed the 126 MB L2; no flush needed" % (B * N / 1e9),
                       "mean_acting_units_per_view": units_mean, "threads_per_game": L.mrts_threads_per_game(env._h), "smem_per_game": L.mrts_smem_bytes(env._h),
                       "staged_units_per_game": L.mrts_staged_units(env._h), "unit_capacity": L.mrts_unit_capacity(env._h),
                       "kernels_per_step": L.mrts_launches_per_step(env._h), "pipeline": {1: "direct", 2: "split"}.get(L.mrts_pipeline(env._h)), "chunks": L.mrts_chunks(env._h),
                       "rules_kernels": kernel_info(L, env._h)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": NCU_TRAFFIC.get((args.workload, args.dtype)),
                         "kernel": "mrts_expand_kernel (bulk-store half of the step; mrts_rules_warp_kernel feeds it)" if L.mrts_pipeline(env._h) == 2 else ("mrts_rules_warp_kernel" if L.mrts_threads_per_game(env._h) == 32 else "mrts_rules_block_kernel"),
                         "kernel_ms_note": "CUDA events around each step on the launching stream (includes the tiny launch-order / overflow kernels of bot slabs)", "bytes_per_env_step": B, "env_steps_per_launch": N, "kernel_ms": kern_ms, "peak_source": peak_src},
            "e2e": e2e, "gpu_launches": args.steps * L.mrts_launches_per_step(env._h), "clocks": clk,
        }
        if cpu_v is not None:
            line["cpu_baseline"] = {"value": cpu_v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                                    "sample": f"{args.cpu_envs} envs of the workload, {cpu_k} steps ({cpu_spent:.1f} s timed inside a {args.cpu_budget:.0f} s leg), get_masks + step timed, action sampling excluded"}
        emit(line)
    env.close()
    if dist is not None:
        dist.destroy_process_group()


def kernel_info(L, h):
    import ctypes as C

    out = {}
    for name, k in (("light", 0), ("heavy", 1)):
        r, g, sm = C.c_int(), C.c_int(), C.c_int()
        L.mrts_kernel_info(h, k, C.byref(r), C.byref(g), C.byref(sm))
        out[name] = {"regs": r.value, "games_per_block": g.value, "smem_per_game": sm.value}
    return out


def e2e_run(args, env_dev, wl, rank, dist, dev):
    """The call a user of the reference makes (ppo_gridnet.py:459-480 loop shape): get_action_mask()
    -> NumPy, policy output -> NumPy, step(actions) -> NumPy obs / reward / done, all int32 / float64
    host arrays as in the reference.  Every step moves the actions host->device and obs + mask + raw
    rewards + dones device->host inside the timed region; the env's own traffic is bit-packed
    (mrts_step_host_slot).  The policy stand-in between two env calls is not timed in `value` (the reference
    arm does not time its action sampling either); `value_incl_policy_copy` is the whole loop."""
    import torch

    from microrts_py_b200 import microrts_ai
    from microrts_py_b200.envs import MicroRTSGridModeVecEnv

    n = min(wl["n"], args.e2e_envs)
    nsp = int(n * (wl["nsp"] / wl["n"])) // 2 * 2
    nbot = n - nsp
    ai2s = [getattr(microrts_ai, wl["bot_names"][i * wl["nbot"] // nbot]) for i in range(nbot)] if nbot else []
    np_dtype = np.int32 if args.dtype == "int32" else np.uint8
    env = MicroRTSGridModeVecEnv(nsp, nbot, partial_obs=wl["partial"], max_steps=wl["max_steps"], ai2s=ai2s, map_paths=[wl["map"]],
                                 device=dev.index, dtype=np_dtype, host_transport=True, seed=77 + rank)
    a_host = torch.zeros((n, env.height * env.width, 7), dtype=torch.int32).pin_memory()
    env.reset()
    mask = env.get_action_mask()
    steps = max(3, args.steps // 4)

    # Like the reference arm (cpu_run: "get_masks + step timed, action sampling excluded") the timed region
    # is the env's own API: step(host actions) + get_action_mask(), host arrays in and out, every
    # host<->device copy of the env inside.  The stand-in policy (device sampler + the 58.7 MB int32
    # `action.cpu().numpy()` copy at 16x16) is the consumer's work: it runs between the timed regions and
    # the loop's wall time including it is reported beside as `value_incl_policy_copy`.
    def one():
        a_dev = env.sample_actions()           # policy stand-in, on the device
        a_host.copy_(a_dev, non_blocking=True)  # action.cpu().numpy()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        obs, r, d, infos = env.step(a_host.numpy())
        m = env.get_action_mask()
        return obs, m, r, d, time.perf_counter() - t0

    for _ in range(3):
        one()
    torch.cuda.synchronize(dev)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    dt_env = 0.0
    for _ in range(steps):
        obs, m, r, d, dt1 = one()
        dt_env += dt1
    torch.cuda.synchronize(dev)
    dt_loop = time.perf_counter() - t0
    t = torch.tensor([dt_env, dt_loop], dtype=torch.float64, device=dev)
    world = 1
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        world = dist.get_world_size()
    dt, dt_loop = float(t[0].item()), float(t[1].item())
    hw = env.height * env.width
    h2d = n * hw * 2                                            # packed action rows
    d2h = n * hw * 16 + n * 6 * 8 + n                           # obs + mask bits, rewards, dones
    host_threads = env._lib.mrts_host_threads(env._h)
    env.close()
    return {"value": n * world * steps / dt, "unit": "env-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "envs_per_gpu": n, "steps": steps,
            "ms_per_step": 1e3 * dt / steps,
            "value_incl_policy_copy": n * world * steps / dt_loop, "policy_copy_d2h_bytes_per_step": a_host.numel() * 4,
            "timed": "step(host actions) + get_action_mask() per step, summed over the steps (max over ranks); the stand-in policy's device sampling and its action-grid copy to the host run between the timed regions, as the reference arm excludes action sampling",
            "api": "MicroRTSGridModeVecEnv(host_transport=True).step/get_action_mask, NumPy int32 in/out, two alternating result buffers; mrts_step_host_slot moves 1 bit per element over PCIe and %d host threads rewrite the cells of the arrays that changed" % host_threads,
            "result_array_bytes_per_step": int(obs.nbytes + m.nbytes)}


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=list(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="envs per GPU (default: the workload's)")
    ap.add_argument("--dtype", default="int32", choices=["int32", "uint8"])
    ap.add_argument("--prewarm", type=int, default=300, help="untimed steps that spread the games over episode phases")
    ap.add_argument("--cpu-envs", type=int, default=1024)
    ap.add_argument("--cpu-budget", type=float, default=25.0, help="wall seconds of the cpu_baseline leg, untimed action sampling included (0 = skip)")
    ap.add_argument("--e2e-envs", type=int, default=8192)
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps == 200 and args.warmup == 20:
            args.steps, args.warmup = 30, 3
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
