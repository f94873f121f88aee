"""TEST INFRASTRUCTURE -- NOT PRODUCT CODE, NOT A FALLBACK.

ctypes wrapper over tests/hostsim/libmrts_hostsim.so: the device rule/encode headers compiled
for the host with a thread group of one (see hostsim.cpp).  Used by the CPU test-suite to check
the kernels' logic against the oracle where no GPU exists.  microrts_py_b200 never imports it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from microrts_py_b200._capi import ConfigC, make_config

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libmrts_hostsim.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-C", _HERE, "libmrts_hostsim.so"], stdout=subprocess.DEVNULL)
        L = C.CDLL(_LIB)
        vp = C.c_void_p
        L.hs_create.argtypes = [C.POINTER(ConfigC), C.POINTER(vp)]
        L.hs_destroy.argtypes = [vp]
        L.hs_reset.argtypes = [vp, vp, vp]
        L.hs_step.argtypes = [vp, vp, vp, vp, vp, vp]
        L.hs_get_masks.argtypes = [vp, vp]
        L.hs_planes.argtypes = [vp]
        L.hs_set_cycle_maps.argtypes = [vp, vp, C.c_int]
        L.hs_unit_capacity.argtypes = [vp]
        L.hs_export_state.argtypes = [vp, C.c_int, vp, vp, C.c_int]
        L.hs_import_state.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp]
        L.hs_can_execute_disagreements.argtypes = [vp, C.c_int, C.POINTER(C.c_int64)]
        L.hs_last_error.restype = C.c_char_p
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class HostSimVecEnv:
    """Same call surface as oracle.oracle_py.OracleVecEnv."""

    def __init__(self, num_selfplay, num_bot, max_steps, partial_obs, templates, ai_ids, none_mode=1, elem_bytes=4, cycle_templates=(), ai1_ids=None):
        self.L = lib()
        cfg, self._keep, cyc = make_config(
            num_selfplay, num_bot, max_steps, partial_obs, templates, ai_ids, none_mode=none_mode, elem_bytes=elem_bytes, cycle_templates=list(cycle_templates), bot_ai1=ai1_ids
        )
        h = C.c_void_p()
        rc = self.L.hs_create(C.byref(cfg), C.byref(h))
        if rc:
            raise ValueError(self.L.hs_last_error().decode())
        self.h = h
        if len(cyc) and self.L.hs_set_cycle_maps(self.h, cyc.ctypes.data, len(cyc)):
            raise ValueError("cycle maps rejected")
        self.N = num_selfplay + num_bot
        self.H, self.W = templates[0].height, templates[0].width
        self.P = self.L.hs_planes(self.h)
        self.dtype = np.int32 if elem_bytes == 4 else np.uint8

    def __del__(self):
        if getattr(self, "h", None):
            self.L.hs_destroy(self.h)
            self.h = None

    def reset(self):
        obs = np.full((self.N, self.H, self.W, self.P), 77, self.dtype)
        self.L.hs_reset(self.h, _p(obs), None)
        return obs

    def get_masks(self):
        m = np.full((self.N, self.H * self.W, 78), 77, self.dtype)
        self.L.hs_get_masks(self.h, _p(m))
        return m, m[:, :, 0].copy()

    def step(self, actions, with_mask=False):
        a = np.ascontiguousarray(np.asarray(actions).reshape(self.N, self.H * self.W, 7), np.int32)
        obs = np.full((self.N, self.H, self.W, self.P), 77, self.dtype)
        rew = np.full((self.N, 6), 77, np.float64)
        done = np.full((self.N,), 77, np.uint8)
        mask = np.full((self.N, self.H * self.W, 78), 77, self.dtype) if with_mask else None
        self.L.hs_step(self.h, _p(a), _p(obs), _p(mask) if with_mask else None, _p(rew), _p(done))
        if with_mask:
            return obs, rew, done, mask
        return obs, rew, done

    def import_state(self, game, hdr5, rows12, issue_order=None):
        rows12 = np.ascontiguousarray(rows12, np.int32).reshape(-1, 12)
        hdr5 = np.ascontiguousarray(hdr5, np.int32)
        order = np.ascontiguousarray(issue_order, np.int32) if issue_order is not None else None
        if self.L.hs_import_state(self.h, game, _p(hdr5), _p(rows12), len(rows12), _p(order) if order is not None else None):
            raise ValueError(self.L.hs_last_error().decode())

    def can_execute_disagreements(self, game):
        """(pairs on which Sim::can_execute and Sim::unit_mask disagree, pairs compared) on the stored state of `game`"""
        n = C.c_int64(0)
        bad = self.L.hs_can_execute_disagreements(self.h, game, C.byref(n))
        return bad, n.value

    def game_state(self, game):
        cap = self.L.hs_unit_capacity(self.h)
        hdr = np.zeros(5, np.int32)
        rows = np.zeros((cap, 12), np.int32)
        n = self.L.hs_export_state(self.h, game, _p(hdr), _p(rows), cap)
        return hdr, rows[:n].copy()
