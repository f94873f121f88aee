"""Unit-level invariants of the rules core (microrts_py_b200/csrc/mrts_core.h), on the CPU build of the same headers.

Sim::can_execute evaluates Unit.canExecuteAction (J/rts/units/Unit.java:531-534: "is the action in getUnitActions?") directly
for one action; Sim::unit_mask enumerates getUnitActions into the 78-bit mask (Unit.java:381-520 via
GameState.getVectorObservation's mask, JNIGridnetClient.getMasks :203-216).  The two must agree on every (unit, action)."""
import numpy as np
import pytest

from microrts_py_b200 import microrts_ai
from tests.backends import make_env
from tests.common import MAP_8, MAP_16, sample_actions

W, LR, P = microrts_ai.workerRushAI, microrts_ai.lightRushAI, microrts_ai.passiveAI


@pytest.mark.parametrize("mp,nsp,ais,steps", [(MAP_16, 2, [W, LR, P, microrts_ai.PORangedRush], 420), (MAP_8, 4, [W, LR], 260)])
def test_can_execute_agrees_with_unit_mask_on_sampled_states(mp, nsp, ais, steps):
    env = make_env("hostsim", num_selfplay_envs=nsp, num_bot_envs=len(ais), max_steps=400, ai2s=ais, map_paths=[mp])
    eng = env.e
    rng = np.random.default_rng(17)
    env.reset()
    n_games = nsp // 2 + len(ais)
    total = 0
    for t in range(steps):
        env.step(sample_actions(env.get_action_mask(), rng, 0.05))
        if t % 7 == 0:
            for g in range(n_games):
                bad, n = eng.can_execute_disagreements(g)
                assert bad == 0, f"step {t}, game {g}: {bad} of {n} (unit, action) pairs disagree"
                total += n
    assert total > 100_000
