"""ctypes binding of libmrts_b200.so (include/mrts_b200.h).

The library is the product: CUDA kernels for sm_100a behind a plain C ABI.  It is built in-tree
(``microrts_py_b200/csrc/libmrts_b200.so``, see ``__graft_entry__.build``) and loaded from there.
There is no fallback of any kind: if the library is missing or no CUDA device is visible the
functions below raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Sequence

import numpy as np

from .maps import MapTemplate

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("MRTS_B200_LIB") or os.path.join(CSRC, "libmrts_b200.so")  # override: tuning builds only

AI_PASSIVE, AI_WORKER_RUSH = 0, 1
NONE_UPSTREAM, NONE_FORK = 0, 1


class MapTemplateC(C.Structure):
    _fields_ = [
        ("width", C.c_int32),
        ("height", C.c_int32),
        ("terrain", C.c_void_p),
        ("resources", C.c_int32 * 2),
        ("n_units", C.c_int32),
        ("units7", C.c_void_p),
    ]


class ConfigC(C.Structure):
    _fields_ = [
        ("num_selfplay_envs", C.c_int32),
        ("num_bot_envs", C.c_int32),
        ("max_steps", C.c_int32),
        ("partial_obs", C.c_int32),
        ("none_mode", C.c_int32),
        ("device", C.c_int32),
        ("n_maps", C.c_int32),
        ("maps", C.POINTER(MapTemplateC)),
        ("env_map", C.c_void_p),
        ("bot_ai", C.c_void_p),
        ("elem_bytes", C.c_int32),
        ("threads_per_game", C.c_int32),
        ("staged_units", C.c_int32),
        ("bot_ai1", C.c_void_p),
        ("pipeline", C.c_int32),
        ("chunks", C.c_int32),
    ]


class PackedIOC(C.Structure):
    _fields_ = [
        ("actions", C.c_void_p),
        ("obs_words", C.c_void_p),
        ("mask_entries", C.c_void_p),
        ("reward", C.c_void_p),
        ("done", C.c_void_p),
        ("next_actions", C.c_void_p),
        ("seed", C.c_uint64),
    ]


PIPELINE_AUTO, PIPELINE_DIRECT, PIPELINE_SPLIT = 0, 1, 2


class StepIOC(C.Structure):
    _fields_ = [
        ("actions", C.c_void_p),
        ("obs", C.c_void_p),
        ("mask", C.c_void_p),
        ("reward", C.c_void_p),
        ("done", C.c_void_p),
        ("next_actions", C.c_void_p),
        ("seed", C.c_uint64),
    ]


def make_config(
    num_selfplay_envs: int,
    num_bot_envs: int,
    max_steps: int,
    partial_obs: bool,
    templates: Sequence[MapTemplate],
    bot_ai: Sequence[int],
    none_mode: int = NONE_FORK,
    device: int = 0,
    elem_bytes: int = 4,
    threads_per_game: int = 0,
    staged_units: int = 0,
    cycle_templates: Sequence[MapTemplate] = None,
    bot_ai1: Sequence[int] = None,
    pipeline: int = 0,
    chunks: int = 0,
):
    """Build an ``mrts_config``.  ``templates`` has one entry per env (as ``map_paths`` after
    broadcasting); identical templates are de-duplicated by object identity / name.
    Returns (cfg, keepalive) -- keep ``keepalive`` referenced while ``cfg`` is in use.  With
    ``cycle_templates`` (the rotation maps of ``mrts_set_cycle_maps``) those templates join the table and
    a third value is returned: their indices in it (int32 array)."""
    n = num_selfplay_envs + num_bot_envs
    if len(templates) != n:
        raise ValueError(f"need one map template per env ({n}), got {len(templates)}")
    if len(bot_ai) != num_bot_envs:
        raise ValueError("need one scripted opponent per bot env")
    uniq: List[MapTemplate] = []
    index = {}
    env_map = np.zeros(n, np.int32)
    for e, t in enumerate(templates):
        key = id(t)
        if key not in index:
            index[key] = len(uniq)
            uniq.append(t)
        env_map[e] = index[key]
    cycle_idx = np.zeros(len(cycle_templates or ()), np.int32)
    for k, t in enumerate(cycle_templates or ()):
        if id(t) not in index:
            index[id(t)] = len(uniq)
            uniq.append(t)
        cycle_idx[k] = index[id(t)]
    arr = (MapTemplateC * len(uniq))()
    keep = [arr, env_map]
    for k, t in enumerate(uniq):
        terr = np.ascontiguousarray(t.terrain, np.uint8)
        units = np.ascontiguousarray(t.units, np.int32).reshape(-1, 7)
        keep += [terr, units]
        arr[k].width = t.width
        arr[k].height = t.height
        arr[k].terrain = terr.ctypes.data
        arr[k].resources[0] = int(t.resources[0])
        arr[k].resources[1] = int(t.resources[1])
        arr[k].n_units = len(units)
        arr[k].units7 = units.ctypes.data
    ai = np.ascontiguousarray(np.asarray(list(bot_ai), np.int32))
    keep.append(ai)
    cfg = ConfigC()
    cfg.num_selfplay_envs = num_selfplay_envs
    cfg.num_bot_envs = num_bot_envs
    cfg.max_steps = max_steps
    cfg.partial_obs = int(bool(partial_obs))
    cfg.none_mode = none_mode
    cfg.device = device
    cfg.n_maps = len(uniq)
    cfg.maps = C.cast(arr, C.POINTER(MapTemplateC))
    cfg.env_map = env_map.ctypes.data
    cfg.bot_ai = ai.ctypes.data if len(ai) else None
    cfg.elem_bytes = elem_bytes
    cfg.threads_per_game = threads_per_game
    cfg.staged_units = staged_units
    cfg.pipeline = pipeline
    cfg.chunks = chunks
    if bot_ai1 is not None:  # bot vs bot (JNIBotClient): scripted player 0 per bot env, -1 = agent
        if len(bot_ai1) != num_bot_envs:
            raise ValueError("need one player-0 entry per bot env")
        a1 = np.ascontiguousarray(np.asarray(list(bot_ai1), np.int32))
        keep.append(a1)
        cfg.bot_ai1 = a1.ctypes.data if len(a1) else None
    if cycle_templates is not None:
        return cfg, keep, cycle_idx
    return cfg, keep


def build_library(force: bool = False) -> str:
    """Compile libmrts_b200.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h", ".cpp"))]
    srcs.append(os.path.join(_HERE, "..", "include", "mrts_b200.h"))
    stale = not os.path.exists(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs if os.path.exists(s))
    if force or stale:
        subprocess.check_call(["make", "-C", CSRC, "-j", "8", "libmrts_b200.so"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None

SYMBOLS = [
    "mrts_create",
    "mrts_destroy",
    "mrts_reset",
    "mrts_step",
    "mrts_get_masks",
    "mrts_step_ex",
    "mrts_sample_actions",
    "mrts_reset_host",
    "mrts_step_host",
    "mrts_reset_host_slot",
    "mrts_step_host_slot",
    "mrts_host_forget_slot",
    "mrts_host_expand_obs_delta",
    "mrts_host_expand_mask_delta",
    "mrts_host_threads",
    "mrts_host_expand_obs",
    "mrts_host_expand_mask",
    "mrts_host_pack_actions",
    "mrts_num_envs",
    "mrts_num_games",
    "mrts_planes",
    "mrts_unit_capacity",
    "mrts_threads_per_game",
    "mrts_smem_bytes",
    "mrts_staged_units",
    "mrts_launches_per_step",
    "mrts_export_state",
    "mrts_import_state",
    "mrts_packed_sizes",
    "mrts_reset_packed",
    "mrts_step_packed",
    "mrts_expand",
    "mrts_expand_static",
    "mrts_host_timing",
    "mrts_host_pack_actions_bitmap",
    "mrts_host_patch_cells",
    "mrts_host_pack_actions_sparse",
    "mrts_pipeline",
    "mrts_chunks",
    "mrts_kernel_info",
    "mrts_debug_issue",
    "mrts_debug_cycle",
    "mrts_debug_timing",
    "mrts_set_cycle_maps",
    "mrts_last_error",
    "mrts_version",
]


def lib():
    """Load libmrts_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). microrts_py_b200 has no CPU fallback."
            )
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.mrts_create.argtypes = [C.POINTER(ConfigC), C.POINTER(vp)]
        L.mrts_destroy.argtypes = [vp]
        L.mrts_destroy.restype = None
        L.mrts_reset.argtypes = [vp, vp, vp, vp]
        L.mrts_step.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.mrts_get_masks.argtypes = [vp, vp, vp]
        L.mrts_step_ex.argtypes = [vp, C.POINTER(StepIOC), vp]
        L.mrts_sample_actions.argtypes = [vp, vp, C.c_uint64, vp]
        for f in ("mrts_num_envs", "mrts_num_games", "mrts_planes", "mrts_unit_capacity", "mrts_threads_per_game", "mrts_smem_bytes", "mrts_staged_units", "mrts_launches_per_step"):
            getattr(L, f).argtypes = [vp]
        L.mrts_export_state.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp]
        L.mrts_import_state.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp, vp]
        L.mrts_packed_sizes.argtypes = [vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.mrts_reset_packed.argtypes = [vp, vp, vp, vp]
        L.mrts_step_packed.argtypes = [vp, C.POINTER(PackedIOC), vp]
        L.mrts_expand.argtypes = [vp, vp, vp, vp, vp, vp]
        L.mrts_expand_static.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp]
        L.mrts_pipeline.argtypes = [vp]
        L.mrts_chunks.argtypes = [vp]
        L.mrts_kernel_info.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.mrts_debug_issue.argtypes = [vp, C.c_int, vp, C.c_int, vp]
        L.mrts_debug_cycle.argtypes = [vp, C.c_int, C.POINTER(C.c_int), vp]
        L.mrts_debug_timing.argtypes = [vp, vp]
        L.mrts_set_cycle_maps.argtypes = [vp, vp, C.c_int]
        L.mrts_reset_host.argtypes = [vp, vp, vp]
        L.mrts_step_host.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp]
        L.mrts_reset_host_slot.argtypes = [vp, vp, vp, C.c_int]
        L.mrts_step_host_slot.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, C.c_int]
        L.mrts_host_forget_slot.argtypes = [vp, C.c_int]
        L.mrts_host_expand_obs_delta.argtypes = [vp, vp, C.c_int64, C.c_int, C.c_int, C.c_int, vp]
        L.mrts_host_expand_mask_delta.argtypes = [vp, vp, C.c_int64, C.c_int, C.c_int, vp]
        L.mrts_host_threads.argtypes = [vp]
        L.mrts_host_timing.argtypes = [vp, vp, C.c_int]
        L.mrts_host_pack_actions_bitmap.argtypes = [vp, C.c_int, vp, C.c_int64, vp]
        L.mrts_host_patch_cells.argtypes = [vp, C.c_int64, C.c_int, C.c_int, vp, vp]
        L.mrts_host_pack_actions_sparse.argtypes = [vp, C.c_int, vp, C.c_int64, vp, C.c_int64, C.POINTER(C.c_int64)]
        L.mrts_host_expand_obs.argtypes = [vp, C.c_int64, C.c_int, C.c_int, C.c_int, vp]
        L.mrts_host_expand_mask.argtypes = [vp, C.c_int64, C.c_int, C.c_int, vp]
        L.mrts_host_pack_actions.argtypes = [vp, C.c_int, vp, C.c_int64, C.c_int, vp]
        L.mrts_last_error.restype = C.c_char_p
        L.mrts_version.restype = C.c_char_p
        _lib = L
    return _lib


def check(rc: int):
    if rc < 0:
        msg = lib().mrts_last_error().decode()
        if rc == -1:
            raise ValueError(msg)
        raise RuntimeError(f"mrts_b200 error {rc}: {msg}")
    return rc
