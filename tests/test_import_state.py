"""mrts_import_state (GameState.fromXML, J/rts/GameState.java:763-915) and, through it, the reference's golden
traces replayed by the PRODUCT kernels: for every entry k of a trace the state recorded there (completed with the
pending assignments, which the oracle's replay of the trace supplies) is imported into a self-play game, the entry's
unit actions are fed through the ordinary action grids of mrts_step, NOOP steps follow until the next entry's time,
and the exported state must equal the recorded snapshot k + 1.  Segments in which the two players' actions collide
in the same frame are skipped: the self-play client issues player 0 before it decodes player 1
(JNIGridnetClientSelfPlay.java:161-169) while the traces were recorded with both issued together.
"hostsim" runs the same pass on the CPU; "cuda" is mrts_rules_warp_kernel via the C ABI."""
import ctypes as C

import numpy as np
import pytest

from microrts_py_b200 import _capi
from microrts_py_b200.maps import load_map
from oracle.oracle_py import NONE_UPSTREAM, OracleGame
from tests.test_oracle_traces import _entry_actions, _entry_units

WANT = ("basesWorkers16x16A", "basesWorkers8x8", "BWDistantResources32x32", "BloodBath", "melee", "4x4", "TwoBasesBarracks16x16", "basesWorkers10x10")
T_MOVE, T_HARVEST, T_RETURN, T_PRODUCE, T_ATTACK = 1, 2, 3, 4, 5
DX, DY = (0, 1, 0, -1), (-1, 0, 1, 0)


def _grid(rows7, W, HW):
    """trace rows (ux, uy, type, param, ax, ay, utype) -> (HW, 7) action grid of UnitAction.fromVectorAction"""
    g = np.zeros((HW, 7), np.int32)
    for ux, uy, t, p, ax, ay, ut in rows7:
        a = g[uy * W + ux]
        a[0] = t
        if t == T_MOVE:
            a[1] = p
        elif t == T_HARVEST:
            a[2] = p
        elif t == T_RETURN:
            a[3] = p
        elif t == T_PRODUCE:
            a[4], a[5] = p, ut
        elif t == T_ATTACK:
            a[6] = (ay - uy + 3) * 7 + (ax - ux + 3)
    return g


def _targets(rows7, W):
    out = set()
    for ux, uy, t, p, ax, ay, ut in rows7:
        if t in (T_MOVE, T_PRODUCE):
            out.add((ux + DX[p]) + (uy + DY[p]) * W)
    return out


class _Device:
    def __init__(self, tmpl):
        self.L = _capi.lib()
        cfg, self.keep = _capi.make_config(2, 0, 10 ** 6, False, [tmpl, tmpl], [], none_mode=_capi.NONE_UPSTREAM)
        self.h = C.c_void_p()
        _capi.check(self.L.mrts_create(C.byref(cfg), C.byref(self.h)))
        self.cap = self.L.mrts_unit_capacity(self.h)
        import torch

        self.torch = torch
        self.act = torch.zeros((2, tmpl.width * tmpl.height, 7), dtype=torch.int32, device="cuda")

    def import_state(self, hdr5, rows12, order):
        vp = C.c_void_p
        rows12 = np.ascontiguousarray(rows12, np.int32)
        order = np.ascontiguousarray(order, np.int32)
        _capi.check(self.L.mrts_import_state(self.h, 0, hdr5.ctypes.data_as(vp), rows12.ctypes.data_as(vp), len(rows12), order.ctypes.data_as(vp), None))

    def step(self, grids):
        self.act.copy_(self.torch.as_tensor(grids))
        _capi.check(self.L.mrts_step(self.h, self.act.data_ptr(), None, None, None, None, None))

    def state(self):
        hdr, rows = np.zeros(5, np.int32), np.zeros((self.cap, 12), np.int32)
        n = _capi.check(self.L.mrts_export_state(self.h, 0, hdr.ctypes.data_as(C.c_void_p), rows.ctypes.data_as(C.c_void_p), self.cap, None))
        return hdr, rows[:n]

    def close(self):
        self.L.mrts_destroy(self.h)


class _Host:
    def __init__(self, tmpl):
        from tests.hostsim import HostSimVecEnv

        self.e = HostSimVecEnv(2, 0, 10 ** 6, False, [tmpl, tmpl], [], none_mode=0)

    def import_state(self, hdr5, rows12, order):
        self.e.import_state(0, hdr5, rows12, order)

    def step(self, grids):
        self.e.step(grids)

    def state(self):
        return self.e.game_state(0)

    def close(self):
        pass


def _replay(golden_traces, make, min_snapshots, max_traces):
    picked = [t for t in golden_traces if any(w in t["map"] for w in WANT)][:max_traces]
    checked = skipped = 0
    for tr in picked:
        tmpl = load_map(tr["map"])
        W, HW = tmpl.width, tmpl.width * tmpl.height
        eng = make(tmpl)
        orc = OracleGame(tmpl, NONE_UPSTREAM)
        entries = tr["entries"]
        noop = np.zeros((2, HW, 7), np.int32)
        for k, e in enumerate(entries[:-1]):
            while orc.time < e[0]:
                orc.cycle()
            per = _entry_actions(e)
            nxt = entries[k + 1]
            collide = bool(_targets(per[0], W) & _targets(per[1], W))
            nu = _entry_units(nxt)
            over = len(set(nu[nu[:, 1] >= 0, 1])) <= 1  # the game ends inside this segment: the vec client restarts it at once
            if not collide and not over and (per[0] or per[1]):
                hdr3, rows = orc.state()  # the recorded snapshot + the assignments pending at it
                np.testing.assert_array_equal(rows[:, :6], _entry_units(e))
                eng.import_state(np.array([hdr3[0], hdr3[1], hdr3[2], 0, len(rows)], np.int32), rows, orc.issue_order())
                eng.step(np.stack([_grid(per[0], W, HW), _grid(per[1], W, HW)]))
                for _ in range(e[0] + 1, nxt[0]):
                    eng.step(noop)
                hdr, got = eng.state()
                where = f"{tr['map']} {tr['ai']} entry {k} -> {k + 1}"
                assert (hdr[0], hdr[1], hdr[2]) == (nxt[0], nxt[1], nxt[2]), where
                np.testing.assert_array_equal(got[:, :6], _entry_units(nxt), err_msg=where)
                checked += 1
            else:
                skipped += collide
            if per[0] or per[1]:
                orc.issue_safe(per[0])
                orc.issue_safe(per[1])
        eng.close()
    assert checked >= min_snapshots, (checked, skipped)
    return checked, skipped


def test_golden_snapshots_through_the_pass_hostsim(golden_traces):
    _replay(golden_traces, _Host, 300, 12)


@pytest.mark.gpu
def test_golden_snapshots_through_the_product_kernels(golden_traces):
    checked, skipped = _replay(golden_traces, _Device, 1000, 10 ** 6)
    print(f"{checked} golden snapshots imported and stepped on the device, {skipped} collision segments skipped")


@pytest.mark.gpu
def test_import_export_round_trip_and_argument_checks():
    tmpl = load_map("maps/16x16/basesWorkers16x16A.xml")
    d = _Device(tmpl)
    hdr, rows = d.state()
    rows = rows.copy()
    rows[0, 4] = 17  # a resource pile
    hdr2 = np.array([123, 9, 4, 55, len(rows)], np.int32)
    d.import_state(hdr2, rows, np.arange(len(rows), dtype=np.int32))
    h, r = d.state()
    np.testing.assert_array_equal(h, hdr2)
    np.testing.assert_array_equal(r, rows)
    bad = rows.copy()
    bad[1, 2] = 99
    with pytest.raises(ValueError):
        d.import_state(hdr2, bad, np.arange(len(rows), dtype=np.int32))
    d.close()
