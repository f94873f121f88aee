"""Pins the oracle's dynamics and bot machinery against the reference's 280 golden traces
(gym_microrts/microrts/data/traces, converted by tests/golden/make_golden.py).

Design mirrored from gym_microrts/microrts/test/microrts/TestTracesIntegrity.java:72-126,
strengthened from "actions are legal" to full state equality at every trace entry.
"""
import numpy as np
import pytest

from microrts_py_b200.maps import load_map
from oracle.oracle_py import AI_LIGHT_RUSH, NONE_FORK, NONE_UPSTREAM, OracleBot, OracleGame


def _entry_units(e):
    return np.asarray(e[3], np.int32).reshape(-1, 6)


def _entry_actions(e):
    """rows (ux, uy, type, param, ax, ay, utype) split by owner, trace order kept."""
    units = _entry_units(e)
    acts = np.asarray(e[4], np.int32).reshape(-1, 6)
    per = {0: [], 1: []}
    for a in acts:
        u = units[a[0]]
        per[int(u[1])].append([u[2], u[3], a[1], a[2], a[3], a[4], a[5]])
    return per


def _assert_state(game, e, where):
    hdr, rows = game.state()
    assert hdr[0] == e[0], where
    assert (hdr[1], hdr[2]) == (e[1], e[2]), f"{where}: player resources"
    np.testing.assert_array_equal(rows[:, :6], _entry_units(e), err_msg=where)


@pytest.mark.parametrize("none_mode", [NONE_UPSTREAM, NONE_FORK])
def test_replay_all_traces_state_equality(golden_traces, none_mode):
    """issueSafe + cycle reproduce every recorded state (unit list order included)."""
    assert len(golden_traces) == 280
    for tr in golden_traces:
        game = OracleGame(load_map(tr["map"]), none_mode)
        for k, e in enumerate(tr["entries"]):
            while game.time < e[0]:
                game.cycle()
            _assert_state(game, e, f"{tr['map']} {tr['ai']} entry {k}")
            per = _entry_actions(e)
            if per[0] or per[1]:
                game.issue_safe(per[0])
                game.issue_safe(per[1])


def test_lightrush_regenerates_traces_action_for_action(golden_traces):
    """LightRush vs LightRush (GenerateTestTraces.java:102-134 loop) under upstream NONE
    semantics reproduces the 140 LightRush traces: same decision times, states and actions.
    This pins AbstractionLayerAI.translateActions, Train/Build/Harvest/Attack and the A*
    tie-breaking that WorkerRush shares."""
    lr = [t for t in golden_traces if t["ai"] == "LightRush"]
    assert len(lr) == 140
    for tr in lr:
        game = OracleGame(load_map(tr["map"]), NONE_UPSTREAM)
        bots = [OracleBot(AI_LIGHT_RUSH), OracleBot(AI_LIGHT_RUSH)]
        entries = tr["entries"]
        where = tr["map"]
        _assert_state(game, entries[0], where + " initial")
        k = 1
        gameover = False
        while True:
            pa = [game.bot_action(bots[p], p) for p in (0, 1)]
            if len(pa[0]) or len(pa[1]):
                e = entries[k]
                _assert_state(game, e, f"{where} entry {k}")
                per = _entry_actions(e)
                for p in (0, 1):
                    np.testing.assert_array_equal(pa[p], np.asarray(per[p], np.int32).reshape(-1, 7), err_msg=f"{where} entry {k} player {p}")
                k += 1
            game.issue_safe(pa[0])
            game.issue_safe(pa[1])
            gameover = game.cycle()
            if gameover or game.time >= 500:
                break
        _assert_state(game, entries[k], where + " final")
        assert k == len(entries) - 1, where
