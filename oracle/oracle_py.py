"""TEST INFRASTRUCTURE -- NOT PRODUCT CODE.

ctypes wrapper over oracle/libmrts_oracle.so (the CPU restatement of the reference's Java
path).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmrts_oracle.so")
_lib = None

NONE_UPSTREAM, NONE_FORK = 0, 1
AI_PASSIVE, AI_WORKER_RUSH, AI_LIGHT_RUSH, AI_PO_WORKER_RUSH, AI_PO_LIGHT_RUSH, AI_PO_HEAVY_RUSH, AI_PO_RANGED_RUSH = 0, 1, 2, 3, 4, 5, 6


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".hpp"))]
    if force or not os.path.exists(_LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libmrts_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, ip = C.c_void_p, C.POINTER(C.c_int32)
        L.orc_vec_create.restype = vp
        L.orc_vec_create.argtypes = [C.c_int] * 6
        L.orc_vec_add_env.argtypes = [vp, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_vec_init.argtypes = [vp]
        L.orc_vec_set_threads.argtypes = [vp, C.c_int]
        L.orc_vec_sample_actions.argtypes = [vp, C.c_void_p, C.c_uint64, C.c_void_p]
        L.orc_vec_set_ai1.argtypes = [vp, C.c_int, C.c_int]
        L.orc_vec_reset.argtypes = [vp, C.c_void_p]
        L.orc_vec_masks.argtypes = [vp, C.c_void_p, C.c_void_p]
        L.orc_vec_step.argtypes = [vp, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_vec_planes.argtypes = [vp]
        L.orc_vec_swap_map.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_vec_game_state.argtypes = [vp, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_vec_destroy.argtypes = [vp]
        L.orc_game_create.restype = vp
        L.orc_game_create.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_game_destroy.argtypes = [vp]
        L.orc_game_issue_safe.argtypes = [vp, C.c_int, C.c_void_p]
        L.orc_game_cycle.argtypes = [vp]
        L.orc_game_time.argtypes = [vp]
        L.orc_game_winner.argtypes = [vp]
        L.orc_game_state.argtypes = [vp, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_game_issue_order.argtypes = [vp, C.c_void_p, C.c_int]
        L.orc_game_masks.argtypes = [vp, C.c_int, C.c_void_p]
        L.orc_bot_create.restype = vp
        L.orc_bot_create.argtypes = [C.c_int]
        L.orc_bot_destroy.argtypes = [vp]
        L.orc_bot_action.argtypes = [vp, vp, C.c_int, C.c_void_p, C.c_int]
        L.orc_version.restype = C.c_char_p
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class OracleGame:
    """One game: low-level issueSafe / cycle access for trace replay and bot regeneration."""

    def __init__(self, tmpl, none_mode=NONE_FORK):
        self.L = lib()
        self.W, self.H = tmpl.width, tmpl.height
        terr = np.ascontiguousarray(tmpl.terrain, np.uint8)
        units = np.ascontiguousarray(tmpl.units, np.int32)
        self.h = self.L.orc_game_create(self.W, self.H, _p(terr), tmpl.resources[0], tmpl.resources[1], len(units), _p(units), none_mode)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_game_destroy(self.h)
            self.h = None

    def issue_safe(self, rows):
        rows = np.ascontiguousarray(np.asarray(rows, np.int32).reshape(-1, 7))
        return self.L.orc_game_issue_safe(self.h, len(rows), _p(rows))

    def cycle(self):
        return bool(self.L.orc_game_cycle(self.h))

    @property
    def time(self):
        return self.L.orc_game_time(self.h)

    def winner(self):
        return self.L.orc_game_winner(self.h)

    def state(self):
        cap = self.W * self.H
        hdr = np.zeros(3, np.int32)
        rows = np.zeros((cap, 12), np.int32)
        n = self.L.orc_game_state(self.h, _p(hdr), _p(rows), cap)
        return hdr, rows[:n].copy()

    def issue_order(self):
        """position of every unit's assignment in the unitActions map (issue order), -1 = idle, in unit order"""
        out = np.zeros(self.W * self.H, np.int32)
        n = self.L.orc_game_issue_order(self.h, _p(out), len(out))
        return out[:n].copy()

    def masks(self, player):
        m = np.zeros((self.H, self.W, 79), np.int32)
        self.L.orc_game_masks(self.h, player, _p(m))
        return m

    def bot_action(self, bot, player):
        rows = np.zeros((self.W * self.H, 7), np.int32)
        n = self.L.orc_bot_action(self.h, bot.h, player, _p(rows), len(rows))
        return rows[:n].copy()


class OracleBot:
    def __init__(self, ai_id):
        self.L = lib()
        self.h = self.L.orc_bot_create(ai_id)
        assert self.h

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_bot_destroy(self.h)
            self.h = None


class OracleVecEnv:
    """JNIGridnetVecClient-level oracle: raw int32 obs, (N,HW,78) masks, (N,6) rewards, (N,) done."""

    def __init__(self, num_selfplay, num_bot, max_steps, partial_obs, templates, ai_ids, none_mode=NONE_FORK, threads=1, ai1_ids=None):
        self.L = lib()
        self.N = num_selfplay + num_bot
        assert len(templates) == self.N and len(ai_ids) == num_bot
        self.h = self.L.orc_vec_create(num_selfplay, num_bot, max_steps, int(partial_obs), none_mode, threads)
        for e, t in enumerate(templates):
            terr = np.ascontiguousarray(t.terrain, np.uint8)
            units = np.ascontiguousarray(t.units, np.int32)
            ai = ai_ids[e - num_selfplay] if e >= num_selfplay else -1
            self.L.orc_vec_add_env(self.h, t.width, t.height, _p(terr), t.resources[0], t.resources[1], len(units), _p(units), ai)
        for b, a1 in enumerate(ai1_ids or ()):  # JNIBotClient games (MicroRTSBotVecEnv): scripted player 0
            self.L.orc_vec_set_ai1(self.h, b, a1)
        self.L.orc_vec_init(self.h)
        self.H, self.W = templates[0].height, templates[0].width
        self.P = self.L.orc_vec_planes(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_vec_destroy(self.h)
            self.h = None

    def reset(self):
        obs = np.zeros((self.N, self.H, self.W, self.P), np.int32)
        self.L.orc_vec_reset(self.h, _p(obs))
        return obs

    def get_masks(self, out=None):
        """(mask (N,HW,78), source_unit_mask (N,HW)); `out` = that pair of arrays to reuse (the Java client reuses its own)."""
        if out is None:
            m = np.zeros((self.N, self.H * self.W, 78), np.int32)
            s = np.zeros((self.N, self.H * self.W), np.int32)
        else:
            m, s = out
        self.L.orc_vec_masks(self.h, _p(m), _p(s))
        return m, s

    def set_threads(self, threads):
        self.L.orc_vec_set_threads(self.h, int(threads))

    def sample_actions(self, mask, seed, out):
        """bench.py's policy stand-in for the CPU arms: masked-uniform actions for `mask` into `out` (N,HW,7) int32."""
        self.L.orc_vec_sample_actions(self.h, _p(mask), C.c_uint64(seed), _p(out))
        return out

    def step(self, actions, out=None):
        a = np.ascontiguousarray(np.asarray(actions).reshape(self.N, self.H * self.W, 7), np.int32)
        if out is None:
            obs = np.zeros((self.N, self.H, self.W, self.P), np.int32)
            rew = np.zeros((self.N, 6), np.float64)
            done = np.zeros((self.N,), np.uint8)
        else:
            obs, rew, done = out
        self.L.orc_vec_step(self.h, _p(a), _p(obs), _p(rew), _p(done))
        return obs, rew, done

    def cycle_maps_after_step(self, obs, done, num_bot_envs, next_map):
        """The map-rotation loop of MicroRTSGridModeVecEnv.step_wait (vec_env.py:1038-1056), index tests
        as written there: `done_idx < num_bot_envs` picks the bot branch.  `next_map` is the
        itertools.cycle over the rotation templates; `obs` is patched in place.  Only all-bot or all-self-play
        envs: with both kinds the reference's index test addresses the wrong clients and obs rows (Java orders
        self-play envs first, JNIGridnetVecClient.java:117-124), which nothing could rely on."""
        assert num_bot_envs in (0, self.N)
        for done_idx, d in enumerate(done):
            if done_idx < num_bot_envs:
                if d:
                    self._swap(self.N - num_bot_envs + done_idx, next(next_map), obs)  # clients[done_idx]
            elif d and done_idx % 2 == 0:
                self._swap(2 * ((done_idx - num_bot_envs) // 2), next(next_map), obs)  # selfPlayClients[(done_idx - num_bot_envs) // 2]
        return obs

    def _swap(self, env, t, obs):
        assert (t.height, t.width) == (self.H, self.W) and obs.flags.c_contiguous and obs.dtype == np.int32
        terr = np.ascontiguousarray(t.terrain, np.uint8)
        units = np.ascontiguousarray(t.units, np.int32)
        self.L.orc_vec_swap_map(self.h, env, t.width, t.height, _p(terr), t.resources[0], t.resources[1], len(units), _p(units), _p(obs))

    def game_state(self, env):
        cap = self.W * self.H
        hdr = np.zeros(3, np.int32)
        rows = np.zeros((cap, 12), np.int32)
        n = self.L.orc_vec_game_state(self.h, env, _p(hdr), _p(rows), cap)
        return hdr, rows[:n].copy()
