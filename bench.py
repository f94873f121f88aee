#!/usr/bin/env python
"""Throughput of the vectorised MicroRTS environment step (BASELINE.json metric: env-steps/s
including observation + mask + reward, and achieved HBM GB/s against the measured peak).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--envs E] [--dtype int32|uint8] [--packed]
  python bench.py --impl reference ...      # the CPU restatement of the reference path (oracle/) on host cores

One "step" = one pass of the hot path over one slab of environments: mrts_step_ex =
JNIGridnetVecClient.gameStep fused with getMasks, writing obs (N,H,W,P), mask (N,HW,78), raw
rewards (N,6) and dones (N,) to HBM.  One env-step = one player view advanced one frame.
`value` times that with actions already resident in HBM (drawn on the device by the fused
masked-uniform sampler); `e2e` times the reference-facing class with HOST buffers: actions
host->device, observation / mask / rewards / dones device->host every step.
Multi-GPU: one process per GPU (torchrun), independent slabs, no collective in the step path
(weak scaling); NCCL is used only for the start/stop barrier and the max-over-ranks time.

Every arm (device, e2e, cpu_baseline, --impl reference) runs the same workload from the same kind of state: a slab
pre-warmed for `--prewarm` steps of masked-uniform play; the e2e arm and both CPU legs run the same number of envs.
The lit_* workloads are the literal BASELINE shapes (N = 2 / N = 24) through the default NumPy API: a latency measurement.
"""
import argparse
import ctypes as C
import hashlib
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
    "cfg5_bloodbath64": (MAP_64, 29, 4096, 1.0, (), False, 2000, "64x64 (4)BloodBath.scmB self-play, 4,096 envs per GPU"),
    "cfg5_dense64": (MAP_64, 29, 4096, 0.0, ("lightRushAI",), False, 6000, "64x64 (4)BloodBath.scmB, dense state: bot envs vs lightRushAI fast-forwarded 1,500 frames (unit-dense combat)"),
    # the literal BASELINE.json configs: what a user who swaps the import gets (default NumPy API, latency-bound)
    "lit_cfg1": (MAP_16, 29, 2, 1.0, (), False, 2000, "BASELINE configs[0] literally: 2 self-play views, 16x16, default NumPy API"),
    "lit_cfg2": (MAP_16, 29, 24, 0.0, ("passiveAI", "workerRushAI"), False, 2000, "BASELINE configs[1] literally: 24 bot envs (12 passiveAI, 12 workerRushAI), default NumPy API"),
}
DEFAULT_WORKLOAD = "cfg2_rollout16"
PREWARM = {"cfg5_dense64": 1500}
CELLS = {MAP_8: 64, MAP_16: 256, MAP_32: 1024, MAP_64: 4096}

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


def packed_bytes(hw, cap):
    """the packed form: one word per cell + the mask entries (counted at their full stride; ~0.3 KB of them are written)"""
    return hw * 4 + (1 + cap) * 16 + hw * 7 * 4 + 49


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def source_hash():
    """identity of the kernel sources (device code + launch logic) a capture under profiles/ was taken on (the GPU box has no .git)"""
    h = hashlib.sha1()
    d = os.path.join(ROOT, "microrts_py_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h")) and not f.startswith("mrts_host"):  # device code and launch logic; the host codec moves no DRAM byte
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def captured_traffic(workload, dtype):
    """dram__bytes_read.sum + dram__bytes_write.sum per step from the ncu capture tools/ncu_traffic.py wrote; null
    (with the reason) when there is none for this workload or it was taken on other kernel sources."""
    p = os.path.join(ROOT, "profiles", f"traffic_{workload}_{dtype}.json")
    if not os.path.exists(p):
        return None, "no capture (tools/ncu_traffic.py)"
    t = json.load(open(p))
    if t.get("source_hash") != source_hash():
        return None, f"capture is stale: taken on sources {t.get('source_hash')}, these are {source_hash()}"
    return t["dram_bytes_per_step"], {k: t[k] for k in ("commit", "kernels", "dram_read_bytes", "dram_write_bytes", "algorithmic_bytes_per_step") if k in t}


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
    return dict(name=name, map=mp, planes=planes, n=n, nsp=nsp, nbot=nbot, bot_names=bot_names, partial=partial, max_steps=max_steps, desc=desc)


def scaled(wl, n):
    """the workload at n envs (same mix)"""
    n = min(n, wl["n"])
    nsp = int(n * (wl["nsp"] / wl["n"])) // 2 * 2
    nbot = n - nsp
    names = [wl["bot_names"][i * wl["nbot"] // nbot] for i in range(nbot)] if nbot else []
    return n, nsp, nbot, names


def cpu_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (oracle/, a C++ restatement of the reference's Java path -- the JVM/JPype
# reference itself cannot run on this image) on the host cores.  The only place besides tests/ and
# smoke() that executes oracle/ code, and only as the thing compared against, never shipped.
# ------------------------------------------------------------------------------------------------
class CpuArm:
    """n envs of the workload on the oracle: pre-warm with masked-uniform play, then time get_masks + step.  The action
    sampling between the two timed calls (the policy stand-in, the oracle's own C sampler) is NOT timed -- as on the GPU arms.
    The result arrays are allocated once and reused, as the Java client reuses its own."""

    AI = {"passiveAI": 0, "workerRushAI": 1, "lightRushAI": 2}

    def __init__(self, wl, n_envs, threads):
        from oracle.oracle_py import OracleVecEnv
        from tests.common import templates_for

        self.n, nsp, nbot, names = scaled(wl, n_envs)
        hw = CELLS[wl["map"]]
        self.env = OracleVecEnv(nsp, nbot, wl["max_steps"], wl["partial"], templates_for([wl["map"]] * self.n), [self.AI[b] for b in names], threads=threads)
        self.obs = self.env.reset()
        self.out = (self.obs, np.zeros((self.n, 6)), np.zeros(self.n, np.uint8))
        self.mask = (np.zeros((self.n, hw, 78), np.int32), np.zeros((self.n, hw), np.int32))
        self.act = np.zeros((self.n, hw, 7), np.int32)
        self.t = 0

    def one(self):
        t0 = time.perf_counter()
        m, _ = self.env.get_masks(self.mask)
        t1 = time.perf_counter()
        self.env.sample_actions(m, 12345 + self.t, self.act)
        t2 = time.perf_counter()
        self.env.step(self.act, self.out)
        t3 = time.perf_counter()
        self.t += 1
        return (t1 - t0) + (t3 - t2)

    def prewarm(self, steps, budget_s):
        t0, k = time.perf_counter(), 0
        while k < steps and time.perf_counter() - t0 < budget_s:
            self.one()
            k += 1
        return k

    def timed(self, threads, min_s, min_steps, max_steps=10 ** 9):
        self.env.set_threads(threads)
        self.one()
        spent, k = 0.0, 0
        while (spent < min_s or k < min_steps) and k < max_steps:
            spent += self.one()
            k += 1
        return self.n * k / spent, k, spent


def cpu_baseline_leg(wl, n_envs, prewarm, budget_s):
    cores = cpu_cores()
    arm = CpuArm(wl, n_envs, cores)
    done = arm.prewarm(prewarm, 0.6 * budget_s)
    v_all, k_all, s_all = arm.timed(cores, 1.0, 3)
    v_one, k_one, s_one = arm.timed(1, 1.0, 2, max_steps=8)
    return {"value": v_all, "unit": "env-steps/s", "cores": cores, "kind": "port", "value_1thread": v_one,
            "sample": f"{arm.n} envs of the workload pre-warmed {done} steps; all {cores} cores: {k_all} steps in {s_all:.2f} s; 1 thread: {k_one} steps in {s_one:.2f} s; "
                      "get_masks + step timed, action sampling excluded",
            "note": "C++ restatement of the reference's Java path (oracle/); the JVM/JPype reference is unavailable on this box"}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = workload_setup(args.workload, args.envs)
    cores = cpu_cores()
    arm = CpuArm(wl, args.cpu_envs or args.e2e_envs or wl["n"], cores)
    pre = arm.prewarm(args.prewarm if args.prewarm >= 0 else PREWARM.get(args.workload, 300), 90.0)
    for _ in range(args.warmup):
        arm.one()
    spent = sum(arm.one() for _ in range(args.steps))
    v = arm.n * args.steps / spent
    sample = f"{arm.n} envs of the workload per step, pre-warmed {pre} steps, {args.steps} timed steps ({spent:.2f} s), get_masks + step timed, action sampling excluded"
    line = {
        "impl": "reference", "metric": "env-steps/sec incl. obs+mask+reward", "value": v, "unit": "env-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * spent / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic (real reference maps, seeded masked-uniform actions)",
        "config": {"workload": args.workload, "description": wl["desc"], "map": wl["map"], "envs_per_step": arm.n, "cells": CELLS[wl["map"]], "prewarm_steps": pre},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample,
                         "note": "C++ restatement of the reference's Java path (oracle/); the JVM/JPype reference is unavailable on this box"},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def kernel_info(L, h):
    import ctypes as C

    out = {}
    for name, k in (("light", 0), ("heavy", 1)):
        r, g, sm = C.c_int(), C.c_int(), C.c_int()
        L.mrts_kernel_info(h, k, C.byref(r), C.byref(g), C.byref(sm))
        out[name] = {"regs": r.value, "games_per_block": g.value, "smem_per_game": sm.value}
    return out


def auto_e2e_envs(workload, dtype):
    """The e2e arm (and the CPU legs beside it) run the workload's own env count when the host can hold the result arrays:
    two slot pairs + the full-write leg's pair + pinned staging ~ 3.5 x (obs + mask) bytes per env and rank."""
    wl = workload_setup(workload, 0)
    if workload.startswith("lit_"):
        return wl["n"]
    per_env = CELLS[wl["map"]] * (wl["planes"] + 78) * (4 if dtype == "int32" else 1)
    lws = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    avail = 64 << 30
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                avail = int(ln.split()[1]) * 1024
    except Exception:  # noqa: BLE001
        pass
    n = wl["n"]
    while n > 2048 and 3.5 * per_env * n * lws > 0.4 * avail:
        n //= 2
    return n


def pin_rank(local):
    """Several ranks on one node: give each its own contiguous block of the cores the job may use (threads that float over every
    core of a two-socket host miss into remote memory, and eight ranks' pools get in one another's way), and size the library's
    host pool to that block.  Opt-in (MRTS_BENCH_PIN=1): the boxes measured are single-socket VMs without SMT, where it changed
    nothing (2 GPUs, cfg2 e2e: 18.3 M pinned, 18.9 M floating); any failure leaves the affinity as it was."""
    lws = int(os.environ.get("LOCAL_WORLD_SIZE", "1"))
    if lws <= 1 or os.environ.get("MRTS_BENCH_PIN", "0") != "1":
        return
    try:
        cores = sorted(os.sched_getaffinity(0))
        k = len(cores) // lws
        if k < 1:
            return
        os.sched_setaffinity(0, set(cores[local * k:(local + 1) * k]))
        os.environ.setdefault("MRTS_HOST_THREADS", str(k))
    except Exception:  # noqa: BLE001
        pass


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
    pin_rank(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    wl = workload_setup(args.workload, args.envs)
    if args.workload.startswith("lit_"):
        return literal_arm(args, wl, rank, local)
    prewarm = args.prewarm if args.prewarm >= 0 else PREWARM.get(args.workload, 300)
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
    cap = L.mrts_unit_capacity(env._h)
    t_dtype = torch.int32 if elem == 4 else torch.uint8
    rew = torch.empty((N, 6), dtype=torch.float64, device=dev)
    done = torch.empty((N,), dtype=torch.uint8, device=dev)
    actions = torch.zeros((N, HW, 7), dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)
    sp = C.c_void_p(stream.cuda_stream)
    if args.packed:  # the rules half alone: one word per cell, one entry per acting unit
        words = torch.empty((N, HW), dtype=torch.int32, device=dev)
        entries = torch.empty((N, 1 + cap, 4), dtype=torch.int32, device=dev)
        io = _capi.PackedIOC()
        io.actions, io.obs_words, io.mask_entries, io.reward, io.done = actions.data_ptr(), words.data_ptr(), entries.data_ptr(), rew.data_ptr(), done.data_ptr()
        io.next_actions, io.seed = actions.data_ptr(), 1000 + rank
        _capi.check(L.mrts_reset_packed(env._h, words.data_ptr(), entries.data_ptr(), sp))

        def step():
            _capi.check(L.mrts_step_packed(env._h, C.byref(io), sp))
    else:
        obs = torch.empty((N, H, Wd, Pn), dtype=t_dtype, device=dev)
        mask = torch.empty((N, HW, 78), dtype=t_dtype, device=dev)
        io = _capi.StepIOC()
        io.actions, io.obs, io.mask, io.reward, io.done = actions.data_ptr(), obs.data_ptr(), mask.data_ptr(), rew.data_ptr(), done.data_ptr()
        io.next_actions, io.seed = actions.data_ptr(), 1000 + rank
        _capi.check(L.mrts_reset(env._h, obs.data_ptr(), mask.data_ptr(), sp))

        def step():
            _capi.check(L.mrts_step_ex(env._h, C.byref(io), sp))

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    _capi.check(L.mrts_sample_actions(env._h, actions.data_ptr(), C.c_uint64(1000 + rank), sp))
    # spread the games over episode phases so that the timed steps see a mix of early / mid / late states
    for _ in range(prewarm):
        step()
    for _ in range(args.warmup):
        step()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record(stream)
    launches = 0
    for k in range(args.steps):
        step()
        launches += L.mrts_launches_per_step(env._h)
        ev[k + 1].record(stream)
    barrier()
    clk = clocks.stop() if rank == 0 else None
    total_ms = ev[0].elapsed_time(ev[-1])
    per_launch = [ev[k].elapsed_time(ev[k + 1]) for k in range(args.steps)]
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    if args.packed:
        units_mean = float(entries[:, 0, 0].sum().item()) / N
        obs_units = None
    else:
        units_mean = float(mask[:, :, 0].sum().item()) / N  # units that may act per view, end state
        obs_units = float((obs[..., 13] == 0).sum().item()) / N  # units on the map per view
    kinfo = kernel_info(L, env._h)
    info = dict(threads_per_game=L.mrts_threads_per_game(env._h), staged=L.mrts_staged_units(env._h), cap=cap, kps=L.mrts_launches_per_step(env._h),
                pipeline={1: "direct", 2: "split"}.get(L.mrts_pipeline(env._h)), chunks=L.mrts_chunks(env._h))
    env.close()
    del env
    if not args.packed:
        del obs, mask
    torch.cuda.empty_cache()

    # ---------------- end to end: the reference-facing class with HOST buffers
    e2e = e2e_run(args, wl, rank, dist, dev, prewarm) if args.e2e_envs > 0 else None

    if rank == 0:
        B = packed_bytes(HW, cap) if args.packed else algorithmic_bytes(HW, Pn, elem)
        peak, peak_src = measured_peak()
        kern_ms = float(np.mean(per_launch))
        achieved = B * N / (kern_ms * 1e-3) / 1e9
        value = N * world * args.steps / (total_ms_max * 1e-3)
        traffic, traffic_src = captured_traffic(args.workload, "packed" if args.packed else args.dtype)
        line = {
            "metric": "env-steps/sec incl. obs+mask+reward", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic (real reference maps; actions drawn on the device by the fused masked-uniform sampler)",
            "config": {"workload": args.workload, "description": wl["desc"], "map": wl["map"], "envs_per_gpu": N, "selfplay_views": nsp, "bot_envs": nbot,
                       "partial_obs": wl["partial"], "max_steps": wl["max_steps"], "prewarm_steps": prewarm,
                       "outputs": "packed (obs words + mask entries)" if args.packed else "obs (N,H,W,P) + mask (N,HW,78) " + args.dtype,
                       "l2": "outputs per step (%.2f GB) exceed the 126 MB L2; no flush needed" % (B * N / 1e9) if B * N > 2.6e8 else "L2 not flushed: %.0f MB of outputs per step" % (B * N / 1e6),
                       "mean_acting_units_per_view": units_mean, "mean_units_per_view": obs_units, "threads_per_game": info["threads_per_game"],
                       "staged_units_per_game": info["staged"], "unit_capacity": info["cap"], "kernels_per_step": info["kps"], "pipeline": info["pipeline"],
                       "chunks": info["chunks"], "rules_kernels": kinfo},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "kernel": "mrts_rules_warp_kernel (rules + run-length bulk stores of obs / mask)" if info["pipeline"] == "direct" else "mrts_rules_warp_kernel -> mrts_expand_kernel",
                         "kernel_ms_note": "CUDA events around each step on the launching stream (all kernels of the step: light + heavy rules, the full-capacity launch)",
                         "bytes_per_env_step": B, "env_steps_per_launch": N, "kernel_ms": kern_ms, "peak_source": peak_src},
            "e2e": e2e, "gpu_launches": launches, "clocks": clk,
        }
        if world == 1 and args.cpu_budget > 0:
            line["cpu_baseline"] = cpu_baseline_leg(wl, args.cpu_envs or args.e2e_envs or 1024, prewarm, args.cpu_budget)
        emit(line)
    if dist is not None:
        dist.destroy_process_group()


def e2e_run(args, wl, rank, dist, dev, prewarm):
    """The call a user of the reference makes (ppo_gridnet.py:459-480 loop shape): get_action_mask()
    -> NumPy, policy output -> NumPy, step(actions) -> NumPy obs / reward / done, all int32 / float64
    host arrays as in the reference.  Every step moves the actions host->device and obs + mask + raw
    rewards + dones device->host inside the timed region.  The policy stand-in between two env calls is
    not timed in `value` (the reference arm does not time its action sampling either);
    `value_incl_policy_copy` is the whole loop.  Three figures: `value` = host_transport with result-buffer
    slots (the library rewrites only the cells that changed), `value_full_write` = the plain host-buffer calls
    that write every element of the arrays every step, `value_default_numpy` = the default mode (fresh arrays,
    the full int32 tensors cross PCIe)."""
    import torch

    from microrts_py_b200 import microrts_ai
    from microrts_py_b200.envs import MicroRTSGridModeVecEnv

    n, nsp, nbot, names = scaled(wl, args.e2e_envs)
    ai2s = [getattr(microrts_ai, b) for b in names]
    np_dtype = np.int32 if args.dtype == "int32" else np.uint8
    steps = max(100, args.steps // 2) if args.e2e_steps <= 0 else args.e2e_steps

    def run(mode_kw, steps, prewarm, n_envs=None):
        k = n if n_envs is None else n_envs
        k_sp = int(k * (nsp / n)) // 2 * 2
        env = MicroRTSGridModeVecEnv(k_sp, k - k_sp, partial_obs=wl["partial"], max_steps=wl["max_steps"], ai2s=ai2s[: k - k_sp] if k < n else ai2s, map_paths=[wl["map"]],
                                     device=dev.index, dtype=np_dtype, seed=77 + rank, **mode_kw)
        a_host = torch.zeros((k, env.height * env.width, 7), dtype=torch.int32).pin_memory()
        env.reset()
        env.get_action_mask()

        def one():
            a_dev = env.sample_actions()           # policy stand-in, on the device
            a_host.copy_(a_dev, non_blocking=True)  # action.cpu().numpy()
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            obs, r, d, infos = env.step(a_host.numpy())
            m = env.get_action_mask()
            return obs, m, time.perf_counter() - t0

        # same state distribution as the device arm: pre-warm through the same API
        for _ in range(prewarm):
            one()
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        dt_env = 0.0
        for _ in range(steps):
            obs, m, dt1 = one()
            dt_env += dt1
        torch.cuda.synchronize(dev)
        dt_loop = time.perf_counter() - t0
        t = torch.tensor([dt_env, dt_loop], dtype=torch.float64, device=dev)
        world = 1
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            world = dist.get_world_size()
        ph = (C.c_double * 8)()
        env._lib.mrts_host_timing(env._h, ph, 0)
        calls = max(ph[7], 1.0)
        phases = {"pack_actions_ms": 1e3 * ph[0] / calls, "h2d_launch_ms": 1e3 * ph[1] / calls, "wait_device_ms": 1e3 * ph[2] / calls, "patch_and_copy_out_ms": 1e3 * ph[3] / calls,
                  "h2d_bytes_per_step": ph[4] / calls, "changed_cells_per_env_step": ph[5] / calls / k, "d2h_bytes_per_step": ph[6] / calls}
        res = dict(phases=phases, v=k * world * steps / float(t[0].item()), v_loop=k * world * steps / float(t[1].item()), ms=1e3 * float(t[0].item()) / steps,
                   hw=env.height * env.width, threads=env._lib.mrts_host_threads(env._h), nbytes=int(obs.nbytes + m.nbytes), act=a_host.numel() * 4)
        env.close()
        return res

    slot = run(dict(host_transport=True), steps, prewarm)
    full = run(dict(host_transport="full"), max(10, steps // 4), min(prewarm, 50))
    per_env = slot["nbytes"] // n
    n_dflt = max(2, min(n, int(4e8 // per_env) // 2 * 2))  # fresh arrays: keep the default-mode leg below ~0.4 GB per call
    dflt = run(dict(), max(5, steps // 10), min(prewarm, 20), n_envs=n_dflt)
    hw = slot["hw"]
    d2h = int(slot["phases"]["d2h_bytes_per_step"]) or n * hw * 16 + n * 6 * 8 + n
    return {"value": slot["v"], "unit": "env-steps/s", "h2d_bytes_per_step": int(slot["phases"]["h2d_bytes_per_step"]) or n * hw * 2, "d2h_bytes_per_step": d2h, "d2h_bytes_per_step_full_write": n * hw * 16 + n * 6 * 8 + n,
            "host_phases": slot["phases"], "envs_per_gpu": n, "steps": steps,
            "prewarm_steps": prewarm, "ms_per_step": slot["ms"], "value_incl_policy_copy": slot["v_loop"], "policy_copy_d2h_bytes_per_step": slot["act"],
            "value_full_write": full["v"], "value_default_numpy": dflt["v"], "default_numpy_envs": n_dflt,
            "timed": "step(host actions) + get_action_mask() per step, summed over the steps (max over ranks); the stand-in policy's device sampling and its action-grid copy to the host run between the timed regions, as the reference arm excludes action sampling",
            "api": "MicroRTSGridModeVecEnv(host_transport=True).step/get_action_mask, NumPy int32 in/out, two alternating result buffers; mrts_step_host_slot: the device compares the step's packed words with the slot's shadow in HBM (mrts_delta_kernel) and ONE copy brings count, rewards, dones, the source-cell bitmap and the changed cells to the host, where %d threads rewrite those cells of the arrays; value_full_write = mrts_step_host (every element written every step); value_default_numpy = the default mode (fresh arrays, full tensors over PCIe)" % slot["threads"],
            "result_array_bytes_per_step": slot["nbytes"]}


def literal_arm(args, wl, rank, local):
    """BASELINE configs[0] / configs[1] literally through the default NumPy API: microseconds per get_action_mask() + step()."""
    from microrts_py_b200 import microrts_ai
    from microrts_py_b200.envs import MicroRTSGridModeVecEnv
    from tests.common import sample_actions

    if rank != 0:
        return
    n, nsp, nbot = wl["n"], wl["nsp"], wl["nbot"]
    env = MicroRTSGridModeVecEnv(nsp, nbot, max_steps=wl["max_steps"], ai2s=[getattr(microrts_ai, b) for b in wl["bot_names"]], map_paths=[wl["map"]], device=local)
    rng = np.random.default_rng(0)
    env.reset()
    lat = []
    launches = 0
    for t in range(300 + args.warmup + args.steps):
        t0 = time.perf_counter()
        m = env.get_action_mask()
        t1 = time.perf_counter()
        a = sample_actions(m, rng)  # policy stand-in, not timed
        t2 = time.perf_counter()
        env.step(a)
        t3 = time.perf_counter()
        if t >= 300 + args.warmup:
            lat.append((t1 - t0) + (t3 - t2))
            launches += env._lib.mrts_launches_per_step(env._h)
    hw = env.height * env.width
    env.close()
    us = 1e6 * float(np.mean(lat))
    cpu = CpuArm(wl, n, 1)
    cpu.prewarm(300, 30.0)
    v_cpu, k_cpu, s_cpu = cpu.timed(1, 0.5, 50)
    B = algorithmic_bytes(hw, wl["planes"], 4)
    peak, peak_src = measured_peak()
    achieved = B * n / (us * 1e-6) / 1e9
    emit({
        "metric": "env-steps/sec incl. obs+mask+reward", "value": n / (us * 1e-6), "unit": "env-steps/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": us / 1e3, "latency_us_per_mask_plus_step": us, "latency_us_p50": 1e6 * float(np.median(lat)), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32", "data": "synthetic (real reference map, seeded masked-uniform actions)",
        "config": {"workload": args.workload, "description": wl["desc"], "map": wl["map"], "envs_per_gpu": n, "prewarm_steps": 300, "api": "default NumPy mode: fresh int32 arrays, get_action_mask() + step() timed on the host"},
        "roofline": {"bound": "latency", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                     "note": "N = %d views: a launch + a host round trip per call; the HBM roofline does not bind here" % n, "bytes_per_env_step": B},
        "e2e": {"value": n / (us * 1e-6), "unit": "env-steps/s", "h2d_bytes_per_step": n * hw * 7 * 4, "d2h_bytes_per_step": n * (B - hw * 7 * 4)},
        "cpu_baseline": {"value": v_cpu, "unit": "env-steps/s", "cores": 1, "kind": "port", "latency_us_per_mask_plus_step": 1e6 * n / v_cpu,
                         "sample": f"the same {n} envs on the oracle, 1 thread, {k_cpu} steps in {s_cpu:.2f} s after 300 pre-warm steps"},
        "gpu_launches": launches,
    })


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
    ap.add_argument("--packed", action="store_true", help="device arm writes the packed form (obs words + mask entries) instead of the tensors")
    ap.add_argument("--prewarm", type=int, default=-1, help="untimed steps that spread the games over episode phases (default 300; 1,500 for the dense 64x64 workload)")
    ap.add_argument("--cpu-envs", type=int, default=0, help="envs of the CPU legs (default: the e2e arm's)")
    ap.add_argument("--cpu-budget", type=float, default=40.0, help="wall seconds of the cpu_baseline leg incl. its pre-warm (0 = skip)")
    ap.add_argument("--e2e-envs", type=int, default=-1, help="envs per GPU of the e2e arm and the CPU legs (0 = skip; default: the workload's own count, halved while the host arrays of all local ranks would take more than 40 %% of the available memory)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed steps of the e2e arm (default max(100, steps / 2))")
    args = ap.parse_args()
    if args.e2e_envs < 0:
        args.e2e_envs = auto_e2e_envs(args.workload, args.dtype)
    if args.impl == "reference":
        if args.steps == 200 and args.warmup == 20:
            args.steps, args.warmup = 30, 3
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
