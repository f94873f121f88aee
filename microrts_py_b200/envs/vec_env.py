"""``MicroRTSGridModeVecEnv`` on B200: the reference's vectorised environment class
(gym_microrts/envs/vec_env.py:38-1101) with the JVM/JPype backend replaced by libmrts_b200.so
(include/mrts_b200.h), hand-written sm_100a CUDA kernels that run the whole step on the device.

Same constructor arguments, attributes, method names, shapes and dtypes as the reference; see
INTEGRATION.md for the call-by-call mapping.  PyTorch is used for device memory, streams and
pinned host buffers only.  There is no CPU fallback: without the CUDA library or a CUDA device
the constructor raises.

Extra keyword-only arguments (all default to the reference's behaviour):

  device        CUDA device ordinal or ``torch.device`` (default: the current device)
  dtype         element type of observations and masks: ``np.int32`` (reference) or ``np.uint8``
                (same values, a quarter of the HBM traffic)
  return_torch  return device tensors instead of NumPy arrays (no host round trip)
  zero_copy     NumPy mode only: return views of two alternating pinned host buffers instead of
                fresh arrays (an array stays valid until the second next call of the same method)
  host_transport NumPy mode only: use the library's host-buffer entry points (mrts_reset_host /
                mrts_step_host): results cross PCIe as one bit per element and are widened into the
                NumPy arrays by the library's thread pool; same values, shapes and dtypes.  Arrays are
                two alternating buffers per output (as with ``zero_copy``), handed out READ-ONLY: the
                library remembers (in HBM) what each buffer holds, a kernel finds the cells that changed and only
                those cross PCIe and are rewritten (mrts_*_host_slot), so the buffers must come back unmodified --
                copy before editing
  ai1s          scripted player 0 per bot env (``microrts_ai`` factories, ``None`` = the agent): bot vs bot as in
                the reference's JNIBotClient; the action rows of those envs are ignored
  fuse_mask     compute the action mask of the resulting state inside ``step`` (one launch,
                the state is staged once); ``get_action_mask`` then only hands it out
  return_packed device tensors in the PACKED form (implies ``return_torch``): ``reset`` / ``step`` return the
                observation words (N, H*W) int32 and ``get_action_mask`` the mask entries (N, 1 + cap, 4) int32 of
                ``microrts_py_b200.packed``; ``expand(words, entries)`` widens them on the device (mrts_expand)
  pipeline      "auto" (default), "direct" (the rules pass writes the tensors itself) or "split" (rules kernels
                write a packed form, an expansion kernel streams it into the tensors with bulk async stores while
                the next chunk's rules run); ``chunks`` = slab chunks of the split pipeline (0 = choose)

``infos`` is always a real ``list`` of ``{"raw_rewards": (6,) float64}`` dicts in the NumPy modes, as in the
reference (vec_env.py:1036); with ``return_torch=True`` it is a lazy list-like (``RawRewardInfos``) that supports
``len``, iteration, integer and slice indexing (``list(infos[:])`` as experiments/ppo_gridnet.py:140 does) and
only reads the device when an entry is looked at.
"""
from __future__ import annotations

import ctypes as C
import json
import os
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _capi
from ..maps import MapTemplate, load_map
from ..microrts_ai import DeviceAI
from ..spaces import Box, Discrete, MultiDiscrete

# ---------------------------------------------------------------------------------------------
# UnitTypeTable() = VERSION_ORIGINAL + MOVE_CONFLICT_RESOLUTION_CANCEL_BOTH (J/rts/units/
# UnitTypeTable.java:74-76,105-290), serialised the way UnitTypeTable.toJSON / UnitType.toJSON do
# (UnitTypeTable.java:373-384, UnitType.java:305-350): this is the ``utt`` dict of the reference.
# ---------------------------------------------------------------------------------------------
_UTT_ROWS = [
    # name, cost, hp, minD, maxD, range, produceT, moveT, attackT, harvestT, returnT, harvestAmt, sight, isRes, isStock, canHarv, canMove, canAtk, produces, producedBy
    ("Resource", 1, 1, 1, 1, 1, 10, 10, 10, 10, 10, 1, 0, True, False, False, False, False, [], []),
    ("Base", 10, 10, 1, 1, 1, 250, 10, 10, 10, 10, 1, 5, False, True, False, False, False, ["Worker"], ["Worker"]),
    ("Barracks", 5, 4, 1, 1, 1, 200, 10, 10, 10, 10, 1, 3, False, False, False, False, False, ["Light", "Heavy", "Ranged"], ["Worker"]),
    ("Worker", 1, 1, 1, 1, 1, 50, 10, 5, 20, 10, 1, 3, False, False, True, True, True, ["Base", "Barracks"], ["Base"]),
    ("Light", 2, 4, 2, 2, 1, 80, 8, 5, 10, 10, 1, 2, False, False, False, True, True, [], ["Barracks"]),
    ("Heavy", 2, 4, 4, 4, 1, 120, 12, 5, 10, 10, 1, 2, False, False, False, True, True, [], ["Barracks"]),
    ("Ranged", 2, 1, 1, 1, 3, 100, 10, 5, 10, 10, 1, 3, False, False, False, True, True, [], ["Barracks"]),
]
_UTT_KEYS = [
    "name", "cost", "hp", "minDamage", "maxDamage", "attackRange", "produceTime", "moveTime", "attackTime", "harvestTime",
    "returnTime", "harvestAmount", "sightRadius", "isResource", "isStockpile", "canHarvest", "canMove", "canAttack", "produces", "producedBy",
]  # fmt: skip


def unit_type_table() -> dict:
    types = []
    for i, row in enumerate(_UTT_ROWS):
        d = {"ID": i}
        d.update(dict(zip(_UTT_KEYS, row)))
        d["produces"] = list(d["produces"])
        d["producedBy"] = list(d["producedBy"])
        types.append(d)
    return {"moveConflictResolutionStrategy": 1, "unitTypes": types}


class RewardFunction:
    """Stand-in for the Java reward objects in ``env.rfs``: callers only use ``str(rf)``, the Java
    simple class name (J/ai/reward/RewardFunctionInterface.java:27-29; experiments/ppo_gridnet.py:151-153)."""

    def __init__(self, name: str):
        self.name = name

    def __str__(self):
        return self.name

    __repr__ = __str__


_RF_NAMES = [
    "WinLossRewardFunction",
    "ResourceGatherRewardFunction",
    "ProduceWorkerRewardFunction",
    "ProduceBuildingRewardFunction",
    "AttackRewardFunction",
    "ProduceCombatUnitRewardFunction",
]  # order of vec_env.py:185-195




def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _huge_empty(shape, dtype):
    """np.empty whose pages the kernel may back with transparent huge pages (asked for before the first touch)."""
    a = np.empty(shape, dtype)
    try:
        lo = (a.ctypes.data + 4095) & ~4095
        hi = (a.ctypes.data + a.nbytes) & ~4095
        if hi - lo >= (1 << 22):
            C.CDLL(None, use_errno=True).madvise(C.c_void_p(lo), C.c_size_t(hi - lo), 14)  # MADV_HUGEPAGE; failure is harmless
    except Exception:  # noqa: BLE001
        pass
    return a


def _np_ptr(a: np.ndarray):
    return C.c_void_p(a.ctypes.data)  # also for arrays handed out read-only


class MicroRTSGridModeVecEnv:
    metadata = {"render.modes": ["human", "rgb_array"], "video.frames_per_second": 150}

    def __init__(
        self,
        num_selfplay_envs,
        num_bot_envs,
        partial_obs=False,
        max_steps=2000,
        render_theme=2,
        frame_skip=0,
        ai2s=[],
        map_paths=["maps/10x10/basesTwoWorkers10x10.xml"],
        reward_shaping=True,
        reward_weight=np.array([0.0, 1.0, 0.0, 0.0, 0.0, 5.0]),
        cycle_maps=[],
        autobuild=False,
        jvm_args=[],
        prior_mode="none",
        reward_prior_weight=0.01,
        prior_advice_freq=1,
        seed=1,
        runs_dir=".",
        graph_ttl_file="graph.ttl",
        graph_triples_file="triples.tsv",
        *,
        device=None,
        dtype=np.int32,
        return_torch=False,
        zero_copy=False,
        fuse_mask=True,
        host_transport=False,
        microrts_path=None,
        threads_per_game=0,
        staged_units=0,
        ai1s=None,
        pipeline="auto",
        chunks=0,
        return_packed=False,
    ):
        self.num_selfplay_envs = num_selfplay_envs
        self.num_bot_envs = num_bot_envs
        self.num_envs = num_selfplay_envs + num_bot_envs
        assert self.num_bot_envs == len(ai2s), "for each environment, a microrts ai should be provided"
        self.partial_obs = partial_obs
        self.max_steps = max_steps
        self.render_theme = render_theme
        self.frame_skip = frame_skip
        self.ai2s = ai2s
        self.map_paths = list(map_paths)
        if len(map_paths) == 1:
            self.map_paths = [map_paths[0] for _ in range(self.num_envs)]
        else:
            assert len(map_paths) == self.num_envs, "if multiple maps are provided, they should be provided for each environment"
        self.reward_shaping = reward_shaping
        self.reward_weight = np.asarray(reward_weight, np.float64)
        if str(prior_mode) != "none":
            raise NotImplementedError("the knowledge-graph prior (prior_mode != 'none') is outside the device hot path")
        # per-episode map rotation (vec_env.py:132-133, 1038-1056): done on the device after every step
        self.cycle_maps = list(cycle_maps)
        if self.cycle_maps and num_selfplay_envs > 0 and num_bot_envs > 0:
            raise NotImplementedError(
                "cycle_maps with both self-play and bot envs: the reference's `done_idx < num_bot_envs` test (vec_env.py:1041) "
                "addresses the wrong clients there; use all-bot or all-self-play envs"
            )
        if num_selfplay_envs % 2:
            raise ValueError("num_selfplay_envs must be even (two player views per self-play game)")
        # maps resolve against an explicit root, $MICRORTS_PATH, or the map pack bundled with the package (maps.py)
        self.microrts_path = microrts_path or os.environ.get("MICRORTS_PATH", "")

        # ---- maps: parsed once, device-resident templates (the reference re-parses per reset)
        cache = {}
        templates: List[MapTemplate] = []
        for p in self.map_paths:
            if p not in cache:
                cache[p] = load_map(p, self.microrts_path)
            templates.append(cache[p])
        cycle_templates = []
        for p in self.cycle_maps:
            if p not in cache:
                cache[p] = load_map(p, self.microrts_path)
            cycle_templates.append(cache[p])
        self.height, self.width = templates[0].height, templates[0].width

        # ---- scripted opponents -> device bot ids
        ai_ids = []
        for ai2 in ai2s:
            a = ai2(None) if callable(ai2) else ai2
            if not isinstance(a, DeviceAI):
                raise NotImplementedError(f"ai2s entry {ai2!r} has no device implementation (use microrts_py_b200.microrts_ai)")
            ai_ids.append(a.ai_id)
        # bot vs bot (JNIBotClient, MicroRTSBotVecEnv): scripted player 0 per bot env; its action rows are ignored
        ai1_ids = None
        if ai1s is not None:
            assert len(ai1s) == num_bot_envs, "for each bot environment, a player-0 ai (or None for the agent) should be provided"
            ai1_ids = []
            for ai1 in ai1s:
                a = ai1(None) if callable(ai1) else ai1
                if a is not None and not isinstance(a, DeviceAI):
                    raise NotImplementedError(f"ai1s entry {ai1!r} has no device implementation (use microrts_py_b200.microrts_ai)")
                ai1_ids.append(-1 if a is None else a.ai_id)

        # ---- device
        if not torch.cuda.is_available():
            raise RuntimeError("microrts_py_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", dev_index)
        self._np_dtype = np.dtype(dtype)
        if self._np_dtype not in (np.dtype(np.int32), np.dtype(np.uint8)):
            raise ValueError("dtype must be np.int32 (reference) or np.uint8")
        self._t_dtype = torch.int32 if self._np_dtype == np.int32 else torch.uint8
        self.return_packed = bool(return_packed)
        self.return_torch = bool(return_torch) or self.return_packed
        self.zero_copy = bool(zero_copy)
        self.fuse_mask = bool(fuse_mask)
        self.host_transport = bool(host_transport) and not self.return_torch
        self._host_slots = host_transport != "full"  # "full": the plain host-buffer calls, every element rewritten every step

        self._lib = _capi.lib()
        cfg, self._keep, cycle_idx = _capi.make_config(
            num_selfplay_envs, num_bot_envs, max_steps, partial_obs, templates, ai_ids,
            none_mode=_capi.NONE_FORK, device=dev_index, elem_bytes=self._np_dtype.itemsize, threads_per_game=threads_per_game, staged_units=staged_units,
            cycle_templates=cycle_templates, bot_ai1=ai1_ids,
            pipeline={"auto": _capi.PIPELINE_AUTO, "direct": _capi.PIPELINE_DIRECT, "split": _capi.PIPELINE_SPLIT}[pipeline], chunks=int(chunks),
        )  # fmt: skip
        h = C.c_void_p()
        _capi.check(self._lib.mrts_create(C.byref(cfg), C.byref(h)))
        self._h = h
        if len(cycle_idx):
            _capi.check(self._lib.mrts_set_cycle_maps(self._h, cycle_idx.ctypes.data, len(cycle_idx)))

        # ---- the reference's computed properties (vec_env.py:229-254)
        self.real_utt = None
        self.utt = unit_type_table()
        self.rfs = [RewardFunction(n) for n in _RF_NAMES]
        self.action_space_dims = [6, 4, 4, 4, 4, len(self.utt["unitTypes"]), 7 * 7]
        self.num_planes = [5, 5, 3, len(self.utt["unitTypes"]) + 1, 6, 2]
        if partial_obs:
            self.num_planes = [5, 5, 3, len(self.utt["unitTypes"]) + 1, 6, 2, 2]
        obs_length = sum(self.num_planes)
        assert obs_length == self._lib.mrts_planes(self._h)
        self.observation_space = Box(low=0.0, high=1.0, shape=(self.height, self.width, obs_length), dtype=self._np_dtype.type)
        self.num_planes_len = len(self.num_planes)
        self.num_planes_prefix_sum = [0]
        for num_plane in self.num_planes:
            self.num_planes_prefix_sum.append(self.num_planes_prefix_sum[-1] + num_plane)
        self.action_space = MultiDiscrete(np.array([self.action_space_dims] * self.height * self.width).flatten())
        self.action_plane_space = MultiDiscrete(self.action_space_dims)
        self.source_unit_idxs = np.tile(np.arange(self.height * self.width), (self.num_envs, 1))
        self.source_unit_idxs = self.source_unit_idxs.reshape((self.source_unit_idxs.shape + (1,)))

        # ---- buffers
        N, HW, P = self.num_envs, self.height * self.width, obs_length
        self._shape_obs = (N, self.height, self.width, P)
        self._shape_mask = (N, HW, 78)
        self._shape_act = (N, HW, 7)
        ow, mw = C.c_int64(), C.c_int64()
        _capi.check(self._lib.mrts_packed_sizes(self._h, C.byref(ow), C.byref(mw)))
        self._shape_words = (N, HW)
        self._shape_entries = (N, mw.value // (4 * N), 4)
        self._d_actions = torch.zeros(self._shape_act, dtype=torch.int32, device=self.device)
        self._d_mask: Optional[torch.Tensor] = None  # mask of the current state when valid
        self._mask_valid = False
        self._pending_actions: Optional[torch.Tensor] = None
        self._h_actions = torch.zeros(self._shape_act, dtype=torch.int32).pin_memory()
        self._h_actions_free = None  # event behind the last host->device copy out of _h_actions
        self._rw_t = torch.as_tensor(self.reward_weight, dtype=torch.float64, device=self.device)
        self._host_ring = {}  # name -> [pinned tensors], cursor
        self.source_unit_mask = None
        self._closed = False
        self._sample_seed = int(seed)
        if self.host_transport:
            # np.empty, not np.zeros: NumPy advises the kernel to back large malloc'ed arrays with huge pages but not
            # calloc'ed ones, and the patches of the slot calls touch a few scattered cache lines per changed cell of
            # arrays of hundreds of MB -- on 4 KB pages every one of them is a TLB miss as well.  (The first call on a slot
            # writes every element.)
            self._hbuf = {
                "obs": [_huge_empty(self._shape_obs, self._np_dtype) for _ in range(2)],
                "mask": [_huge_empty(self._shape_mask, self._np_dtype) for _ in range(2)],
                "raw": [np.zeros((N, 6), np.float64) for _ in range(2)],
                "done": [np.zeros((N,), np.uint8) for _ in range(2)],
            }
            for a in self._hbuf["obs"] + self._hbuf["mask"]:
                a.flags.writeable = False  # the library writes through the raw pointer; see host_transport above
            self._hcur = 0
            self._h_mask_np = None

    # ------------------------------------------------------------------------------ helpers
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _new(self, shape, dtype):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def _to_host(self, name: str, t: torch.Tensor) -> np.ndarray:
        """Device tensor -> NumPy through pinned memory.  zero_copy: a view of one of two
        alternating pinned buffers; otherwise a fresh array (the reference's np.array copy)."""
        ring = self._host_ring.get(name)
        if ring is None:
            k = 2 if self.zero_copy else 1
            ring = [[torch.empty(t.shape, dtype=t.dtype).pin_memory() for _ in range(k)], 0]
            self._host_ring[name] = ring
        bufs, cur = ring
        buf = bufs[cur % len(bufs)]
        ring[1] = cur + 1
        buf.copy_(t, non_blocking=True)
        return buf

    def _finish_host(self, *bufs):
        torch.cuda.current_stream(self.device).synchronize()
        out = []
        for b in bufs:
            a = b.numpy()
            out.append(a if self.zero_copy else a.copy())
        return out

    def _check_open(self):
        if self._closed:
            raise RuntimeError("environment is closed")

    # ------------------------------------------------------------------------------ API
    def reset(self):
        """JNIGridnetVecClient.reset + _encode_obs (vec_env.py:278-282): (N,H,W,P) observations."""
        self._check_open()
        if self.host_transport:
            self._hcur ^= 1
            o, m = self._hbuf["obs"][self._hcur], self._hbuf["mask"][self._hcur]
            _capi.check(self._lib.mrts_reset_host_slot(self._h, _np_ptr(o), _np_ptr(m), self._hcur if self._host_slots else -1))
            self._h_mask_np = m
            return o
        if self.return_packed:
            words, entries = self._new(self._shape_words, torch.int32), self._new(self._shape_entries, torch.int32)
            with torch.cuda.device(self.device):
                _capi.check(self._lib.mrts_reset_packed(self._h, _ptr(words), _ptr(entries), self._stream()))
            self._d_mask, self._mask_valid = entries, True
            return words
        obs = self._new(self._shape_obs, self._t_dtype)
        mask = self._new(self._shape_mask, self._t_dtype) if self.fuse_mask else None
        with torch.cuda.device(self.device):
            _capi.check(self._lib.mrts_reset(self._h, _ptr(obs), _ptr(mask), self._stream()))
        self._d_mask, self._mask_valid = mask, mask is not None
        if self.return_torch:
            return obs
        (o,) = self._finish_host(self._to_host("obs", obs))
        return o

    def get_action_mask(self):
        """vec_env.py:1091-1101: (N, H*W, 78) mask of the current state; also refreshes
        ``source_unit_mask`` (N, H*W)."""
        self._check_open()
        if self.host_transport:
            if self._h_mask_np is None:
                raise RuntimeError("get_action_mask() before reset()")
            self.source_unit_mask = self._h_mask_np[:, :, 0]
            return self._h_mask_np
        if self.return_packed:
            if not self._mask_valid:
                raise RuntimeError("return_packed: get_action_mask() hands out the entries of the last reset() / step()")
            self.source_unit_mask = None
            return self._d_mask
        if not self._mask_valid:
            mask = self._new(self._shape_mask, self._t_dtype)
            with torch.cuda.device(self.device):
                _capi.check(self._lib.mrts_get_masks(self._h, _ptr(mask), self._stream()))
            self._d_mask, self._mask_valid = mask, True
        mask = self._d_mask
        if self.return_torch:
            # a unit that may act always has NOOP legal (Unit.java:519): entry 0 of its row == masks[y][x][0]
            self.source_unit_mask = mask[:, :, 0]
            return mask
        (m,) = self._finish_host(self._to_host("mask", mask))
        self.source_unit_mask = m[:, :, 0]
        return m

    def step_async(self, actions):
        """vec_env.py:968-984.  ``actions``: anything reshapeable to (N, H*W, 7), NumPy or torch
        (host or device), any integer dtype.  The reference forwards only the rows of cells whose
        ``source_unit_mask`` is 1; the device decode applies the same rule to the current state
        (PlayerAction.java:407), so the full grid is uploaded as is."""
        self._check_open()
        if self.host_transport:
            a = actions.cpu().numpy() if isinstance(actions, torch.Tensor) else np.asarray(actions)
            a = a.reshape(self._shape_act)
            if a.dtype not in (np.int32, np.int64) or not a.flags.c_contiguous:
                a = np.ascontiguousarray(a, np.int64)
            self._pending_actions = a
            return
        if isinstance(actions, torch.Tensor):
            a = actions.reshape(self._shape_act)
            if a.device.type == "cuda":
                a = a.to(device=self.device, dtype=torch.int32).contiguous()
            else:
                a = self._upload_actions(lambda: self._h_actions.copy_(a))
        else:
            a = np.asarray(actions).reshape(self._shape_act)
            a = self._upload_actions(lambda: np.copyto(self._h_actions.numpy(), a, casting="unsafe"))
        self._pending_actions = a

    def _upload_actions(self, fill):
        """Host actions -> the one pinned staging buffer -> device.  The copy is asynchronous and (return_torch)
        nothing else synchronises between two steps: wait for the previous copy before the buffer is overwritten."""
        if self._h_actions_free is not None:
            self._h_actions_free.synchronize()
        fill()
        self._d_actions.copy_(self._h_actions, non_blocking=True)
        if self._h_actions_free is None:
            self._h_actions_free = torch.cuda.Event()
        self._h_actions_free.record(torch.cuda.current_stream(self.device))
        return self._d_actions

    def step_wait(self):
        """vec_env.py:1001-1057: (obs, reward @ reward_weight, done[:, 0], infos)."""
        self._check_open()
        if self._pending_actions is None:
            raise RuntimeError("step_wait() without step_async()")
        a = self._pending_actions
        self._pending_actions = None
        N = self.num_envs
        if self.host_transport:
            self._hcur ^= 1
            k = self._hcur
            o, m, reward, d = self._hbuf["obs"][k], self._hbuf["mask"][k], self._hbuf["raw"][k], self._hbuf["done"][k]
            vp = C.c_void_p
            _capi.check(self._lib.mrts_step_host_slot(self._h, a.ctypes.data_as(vp), a.dtype.itemsize, _np_ptr(o), _np_ptr(m),
                                                      reward.ctypes.data_as(vp), d.ctypes.data_as(vp), k if self._host_slots else -1))
            self._h_mask_np = m
            if not self.reward_shaping:
                reward[:, 1:] = 0
            # `reward` is one of two alternating buffers: the infos get their own copy, so that rows a wrapper
            # accumulates over an episode do not change two steps later
            return o, reward @ self.reward_weight, d.astype(np.bool_), RawRewardInfos(reward.copy())
        raw = self._new((N, 6), torch.float64)
        done = self._new((N,), torch.uint8)
        if self.return_packed:
            io = _capi.PackedIOC()
            words, entries = self._new(self._shape_words, torch.int32), self._new(self._shape_entries, torch.int32)
            io.actions, io.obs_words, io.mask_entries, io.reward, io.done = a.data_ptr(), words.data_ptr(), entries.data_ptr(), raw.data_ptr(), done.data_ptr()
            with torch.cuda.device(self.device):
                _capi.check(self._lib.mrts_step_packed(self._h, C.byref(io), self._stream()))
            self._d_mask, self._mask_valid = entries, True
            if not self.reward_shaping:
                raw[:, 1:] = 0
            return words, raw @ self._rw_t, done.bool(), RawRewardInfos(raw)
        obs = self._new(self._shape_obs, self._t_dtype)
        mask = self._new(self._shape_mask, self._t_dtype) if self.fuse_mask else None
        with torch.cuda.device(self.device):
            _capi.check(self._lib.mrts_step(self._h, _ptr(a), _ptr(obs), _ptr(mask), _ptr(raw), _ptr(done), self._stream()))
        self._d_mask, self._mask_valid = mask, mask is not None
        if self.return_torch:
            if not self.reward_shaping:
                raw[:, 1:] = 0
            return obs, raw @ self._rw_t, done.bool(), RawRewardInfos(raw)
        o, r, d = self._finish_host(self._to_host("obs", obs), self._to_host("raw", raw), self._to_host("done", done))
        # zero_copy hands out views of two alternating pinned buffers: fine for obs / mask, but the rows of `infos` are
        # accumulated over an episode by the reference's wrappers, so the (tiny) reward block is always a private copy
        reward = r.copy() if self.zero_copy else r
        if not self.reward_shaping:
            reward[:, 1:] = 0
        infos = [{"raw_rewards": item} for item in reward]
        return o, reward @ self.reward_weight, d.astype(np.bool_), infos

    def step(self, ac):
        self.step_async(ac)
        return self.step_wait()

    def sample_actions(self, seed: Optional[int] = None) -> torch.Tensor:
        """Device-side masked-uniform random policy for the current state (the sampling loop of the
        reference's hello_world.py:27-67): returns the (N, H*W, 7) int32 device tensor."""
        self._check_open()
        with torch.cuda.device(self.device):
            _capi.check(self._lib.mrts_sample_actions(self._h, _ptr(self._d_actions), C.c_uint64(self._sample_seed if seed is None else seed), self._stream()))
        return self._d_actions

    def step_into(self, actions, obs, mask=None, raw_rewards=None, done=None, next_actions=None, seed: Optional[int] = None):
        """gameStep + getMasks of the resulting state written straight into the caller's device tensors --
        e.g. ``step_into(a, buf.obs[t + 1], buf.mask[t + 1], buf.raw[t], buf.done[t])`` for a PPO rollout
        buffer (the consumer loop of experiments/ppo_gridnet.py:459-480 without a host hop or an extra copy).
        ``obs`` (N,H,W,P) and ``mask`` (N,H*W,78) have the env's dtype, ``raw_rewards`` (N,6) float64, ``done``
        (N,) uint8, ``next_actions`` (N,H*W,7) int32 receives a masked-uniform random action grid for the
        new state (mrts_step_ex).  All contiguous, on the env's device; nothing is returned or synchronised."""
        self._check_open()

        def chk(t, shape, dtype, name):
            if t is None:
                return None
            if not isinstance(t, torch.Tensor) or t.device != self.device or t.dtype != dtype or tuple(t.shape) != tuple(shape) or not t.is_contiguous():
                raise ValueError(f"{name}: need a contiguous {dtype} tensor of shape {tuple(shape)} on {self.device}")
            return t

        N = self.num_envs
        a = chk(actions, self._shape_act, torch.int32, "actions")
        if a is None:
            raise ValueError("actions is required")
        io = _capi.StepIOC()
        io.actions = a.data_ptr()
        io.obs = chk(obs, self._shape_obs, self._t_dtype, "obs").data_ptr() if obs is not None else None
        io.mask = chk(mask, self._shape_mask, self._t_dtype, "mask").data_ptr() if mask is not None else None
        io.reward = chk(raw_rewards, (N, 6), torch.float64, "raw_rewards").data_ptr() if raw_rewards is not None else None
        io.done = chk(done, (N,), torch.uint8, "done").data_ptr() if done is not None else None
        io.next_actions = chk(next_actions, self._shape_act, torch.int32, "next_actions").data_ptr() if next_actions is not None else None
        io.seed = self._sample_seed if seed is None else seed
        with torch.cuda.device(self.device):
            _capi.check(self._lib.mrts_step_ex(self._h, C.byref(io), self._stream()))
        self._d_mask, self._mask_valid = (mask, True) if mask is not None else (None, False)

    def expand(self, words=None, entries=None):
        """Packed form -> the reference's tensors on the device (mrts_expand: bulk async stores): returns (obs, mask),
        ``None`` for an input that was not given."""
        self._check_open()
        obs = self._new(self._shape_obs, self._t_dtype) if words is not None else None
        mask = self._new(self._shape_mask, self._t_dtype) if entries is not None else None
        with torch.cuda.device(self.device):
            _capi.check(self._lib.mrts_expand(self._h, _ptr(words), _ptr(entries), _ptr(obs), _ptr(mask), self._stream()))
        return obs, mask

    def getattr_depth_check(self, name, already_found):
        if hasattr(self, name) and already_found:
            return "{0}.{1}".format(type(self).__module__, type(self).__name__)
        return None

    def render(self, mode="human"):
        """The reference renders with Java2D (J/tests/JNIGridnetClient.java:144-161); there is no
        graphics stack on the target machines: unavailable."""
        raise NotImplementedError("render() needs the reference's Java2D path: unavailable on this box")

    def close(self):
        if not self._closed and getattr(self, "_h", None):
            self._lib.mrts_destroy(self._h)
            self._h = None
        self._closed = True

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    # ------------------------------------------------------------------------------ cross-checks
    def export_state(self, game: int):
        """(hdr5, rows12) of one game, see mrts_export_state in include/mrts_b200.h."""
        cap = self._lib.mrts_unit_capacity(self._h)
        hdr = np.zeros(5, np.int32)
        rows = np.zeros((cap, 12), np.int32)
        n = _capi.check(self._lib.mrts_export_state(self._h, game, hdr.ctypes.data_as(C.c_void_p), rows.ctypes.data_as(C.c_void_p), cap, self._stream()))
        return hdr, rows[:n].copy()


class MicroRTSBotVecEnv(MicroRTSGridModeVecEnv):
    """Bot-vs-bot games, rewards and dones only (gym_microrts/envs/vec_env.py:1104-1236 over
    J/tests/JNIBotClient.java:108-134): ``ai1s[i]`` plays player 0 and ``ai2s[i]`` player 1 of env i, both on
    the full state; ``step`` ignores its argument and returns ``np.ones((N, 2))`` as the observation, the
    weighted reward of player 0, done and ``[{"raw_rewards": ...}]`` exactly like the reference."""

    def __init__(self, ai1s=[], ai2s=[], partial_obs=False, max_steps=2000, render_theme=2, map_paths="maps/10x10/basesTwoWorkers10x10.xml",
                 reward_weight=np.array([0.0, 1.0, 0.0, 0.0, 0.0, 5.0]), autobuild=False, jvm_args=[], *, device=None, microrts_path=None):
        assert len(ai1s) == len(ai2s), "for each environment, a microrts ai should be provided"
        paths = [map_paths] if isinstance(map_paths, str) else list(map_paths)
        if len(paths) != 1 and len(paths) != len(ai1s):  # the Java client reads mapPaths[i] per env (JNIGridnetVecClient.java:168)
            raise ValueError("map_paths: one path, or one per environment")
        super().__init__(0, len(ai2s), partial_obs=False, max_steps=max_steps, render_theme=render_theme, ai2s=list(ai2s), map_paths=paths,
                         reward_weight=reward_weight, device=device, microrts_path=microrts_path, ai1s=list(ai1s), fuse_mask=False)
        self.ai1s = ai1s
        self.partial_obs = partial_obs  # JNIBotClient ignores it: both bots see the whole state
        if partial_obs:
            self.num_planes = [5, 5, 3, len(self.utt["unitTypes"]) + 1, 6, 2, 2]
        self.observation_space = Discrete(2)
        self.action_space = Discrete(2)
        N = self.num_envs
        self._raw = torch.empty((N, 6), dtype=torch.float64, device=self.device)
        self._done = torch.empty((N,), dtype=torch.uint8, device=self.device)

    def reset(self):
        self._check_open()
        with torch.cuda.device(self.device):
            _capi.check(self._lib.mrts_reset(self._h, None, None, self._stream()))
        return np.ones((self.num_envs, 2))

    def step_async(self, actions):
        self.actions = actions

    def step_wait(self):
        self._check_open()
        with torch.cuda.device(self.device):
            _capi.check(self._lib.mrts_step(self._h, None, None, None, _ptr(self._raw), _ptr(self._done), self._stream()))
        reward = self._raw.cpu().numpy()
        done = self._done.cpu().numpy().astype(np.bool_)
        infos = [{"raw_rewards": item} for item in reward]
        return np.ones((self.num_envs, 2)), reward @ self.reward_weight, done, infos

    def get_action_mask(self):
        raise NotImplementedError("MicroRTSBotVecEnv has no agent: the reference's inherited get_action_mask() fails in getMasks (no clients)")


class RawRewardInfos(Sequence):
    """``infos`` of the torch and host_transport modes: the reference's list of ``{"raw_rewards": (6,) float64}``
    dicts (vec_env.py:1036) over the step's (N, 6) reward block (a device tensor, or a private NumPy copy).  Behaves like that list for the reference's own
    consumers -- ``len``, iteration, ``infos[i]``, ``infos[:]`` / ``list(infos[:])`` and ``infos[i].copy()``
    (MicroRTSStatsRecorder.step_wait, experiments/ppo_gridnet.py:138-158; SB3's VecMonitor) -- but only moves the
    rewards to the host (one copy for all envs) when an entry is first looked at.  ``.raw`` is the device tensor."""

    def __init__(self, raw):
        self.raw = raw
        self._host = raw if isinstance(raw, np.ndarray) else None

    def _rows(self):
        if self._host is None:
            self._host = self.raw.detach().cpu().numpy()
        return self._host

    def __len__(self):
        return self.raw.shape[0]

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [{"raw_rewards": row} for row in self._rows()[i]]
        return {"raw_rewards": self._rows()[i]}

    def __iter__(self):
        return iter([{"raw_rewards": row} for row in self._rows()])
