"""Trajectory parity against the oracle: identical seeded action traces fed to the oracle and to
the engine under test; observations, masks, raw rewards and dones must be bit-exact at every
step, across terminal / max_steps auto-resets.  "hostsim" (CPU) checks the device headers' logic
in the build container; "cuda" is the parity test proper, through MicroRTSGridModeVecEnv and the
C ABI on the B200.  Cases follow BASELINE.json's five configs at sizes the oracle finishes in
seconds; eps > 0 mixes in mask-ignoring actions to exercise the illegal-action rewrite
(GameState.java:347-354)."""
import numpy as np
import pytest

from microrts_py_b200 import microrts_ai
from tests.backends import make_env
from tests.common import MAP_4, MAP_8, MAP_16, MAP_32, MAP_64, sample_actions

P, W, LR = microrts_ai.passiveAI, microrts_ai.workerRushAI, microrts_ai.lightRushAI
POW, POL = microrts_ai.POWorkerRush, microrts_ai.POLightRush
POH, POR = microrts_ai.POHeavyRush, microrts_ai.PORangedRush
RW = np.array([10.0, 1.0, 1.0, 0.2, 1.0, 4.0])

CASES = {
    # name: (num_selfplay, ai2s, map, partial_obs, max_steps, steps, eps)
    "cfg1_selfplay16": (2, [], MAP_16, False, 300, 700, 0.05),
    "cfg2_bots16": (0, [P, W] * 3, MAP_16, False, 400, 900, 0.05),
    "cfg3_selfplay8": (8, [W, P], MAP_8, False, 200, 600, 0.1),
    "cfg4_partial32": (2, [W], MAP_32, True, 150, 250, 0.05),
    "cfg5_bloodbath64": (2, [W], MAP_64, False, 2000, 120, 0.05),
    "partial8_mixed": (4, [W, P], MAP_8, True, 120, 500, 0.1),
    "tiny4": (2, [P, W], MAP_4, False, 50, 300, 0.2),
    # LightRush: builds a barracks, trains light units, every worker harvests
    "lightrush16": (0, [LR, LR, W, LR], MAP_16, False, 700, 1000, 0.03),
    "lightrush8_partial": (0, [LR, LR, P], MAP_8, True, 400, 600, 0.05),
    "lightrush_barracks16": (0, [LR, LR], "maps/16x16/TwoBasesBarracks16x16.xml", False, 600, 700, 0.02),
    "lightrush_nobase8": (0, [LR, LR], "maps/8x8/noBases8x8.xml", False, 300, 400, 0.05),
    # the partially observable variants: melee units explore (Move to the nearest unobservable cell) while no enemy is in view
    "po_bots16_partial": (0, [POW, POL, POW, POL], MAP_16, True, 500, 800, 0.03),
    "po_bots32_partial": (0, [POW, POL], MAP_32, True, 400, 450, 0.03),
    "po_bots8_full_obs": (0, [POW, POL, W], MAP_8, False, 300, 500, 0.05),
    # POHeavyRush / PORangedRush (microrts_ai.py:43-52): the LightRush bot training heavy / ranged units (range-3 attacks)
    "po_heavy_ranged16_partial": (0, [POH, POR, POH, POR], MAP_16, True, 900, 1100, 0.02),
    "po_heavy_ranged_barracks16": (0, [POH, POR, POR], "maps/16x16/TwoBasesBarracks16x16.xml", False, 700, 800, 0.02),
    "po_heavy_ranged8_partial": (0, [POR, POH], MAP_8, True, 500, 700, 0.03),
}


def _engines():
    return [pytest.param("hostsim", id="hostsim"), pytest.param("cuda", id="cuda", marks=pytest.mark.gpu)]


@pytest.mark.parametrize("engine", _engines())
@pytest.mark.parametrize("case", list(CASES))
def test_trajectory_matches_oracle(engine, case):
    nsp, ais, mp, partial, max_steps, steps, eps = CASES[case]
    if engine == "hostsim" and case in ("cfg5_bloodbath64",):
        steps = 60
    kw = dict(num_selfplay_envs=nsp, num_bot_envs=len(ais), partial_obs=partial, max_steps=max_steps, ai2s=ais, map_paths=[mp], reward_weight=RW)
    ref = make_env("oracle", **kw)
    env = make_env(engine, **kw)
    rng = np.random.default_rng(1234)
    np.testing.assert_array_equal(env.reset(), ref.reset(), err_msg="reset observation")
    n_done = 0
    for t in range(steps):
        m_ref = ref.get_action_mask()
        m = env.get_action_mask()
        np.testing.assert_array_equal(m, m_ref, err_msg=f"{case}: mask at step {t}")
        np.testing.assert_array_equal(env.source_unit_mask, ref.source_unit_mask)
        a = sample_actions(m_ref, rng, eps)
        o, r, d, info = env.step(a)
        o2, r2, d2, info2 = ref.step(a)
        np.testing.assert_array_equal(o, o2, err_msg=f"{case}: observation at step {t}")
        np.testing.assert_array_equal(np.stack([i["raw_rewards"] for i in info]), np.stack([i["raw_rewards"] for i in info2]), err_msg=f"{case}: raw rewards at step {t}")
        np.testing.assert_array_equal(r, r2, err_msg=f"{case}: weighted reward at step {t}")
        np.testing.assert_array_equal(d, d2, err_msg=f"{case}: done at step {t}")
        n_done += int(d.sum())
    if steps > max_steps:
        assert n_done > 0, "the case was meant to cross an auto-reset"
    env.close()


# Per-episode map rotation (gym_microrts/envs/vec_env.py:132-133, 1038-1056): every
# client whose episode ended takes the next map of ONE round-robin list (ascending env order) and is
# reset again.  Short episodes so that several envs end in the same step and the list wraps.
CYCLE_CASES = {
    # name: (num_selfplay, ai2s, start map, rotation, partial_obs, max_steps per env?, steps)
    "bots16": (0, [P, W, W, P, W], MAP_16, ["maps/16x16/basesWorkers16x16B.xml", "maps/16x16/TwoBasesBarracks16x16.xml", "maps/16x16/basesWorkers16x16noResources.xml"], False, 37, 400),
    "selfplay8": (6, [], MAP_8, ["maps/8x8/basesWorkers8x8Obstacle.xml", "maps/8x8/FourBasesWorkers8x8.xml", MAP_8, "maps/8x8/melee8x8Mixed6.xml"], False, 29, 300),
    "bots8_partial": (0, [W, W, P], MAP_8, ["maps/8x8/basesWorkers8x8L.xml", "maps/8x8/basesWorkersBarracks8x8.xml"], True, 41, 300),
}


@pytest.mark.parametrize("engine", _engines())
@pytest.mark.parametrize("case", list(CYCLE_CASES))
def test_cycle_maps_match_the_reference_rotation(engine, case):
    nsp, ais, mp, rotation, partial, max_steps, steps = CYCLE_CASES[case]
    kw = dict(num_selfplay_envs=nsp, num_bot_envs=len(ais), partial_obs=partial, max_steps=max_steps, ai2s=ais, map_paths=[mp], reward_weight=RW, cycle_maps=rotation)
    ref = make_env("oracle", **kw)
    env = make_env(engine, **kw)
    rng = np.random.default_rng(99)
    np.testing.assert_array_equal(env.reset(), ref.reset(), err_msg="reset observation")
    n_done = 0
    for t in range(steps):
        m_ref = ref.get_action_mask()
        np.testing.assert_array_equal(env.get_action_mask(), m_ref, err_msg=f"{case}: mask at step {t}")
        a = sample_actions(m_ref, rng, 0.05)
        o, r, d, info = env.step(a)
        o2, r2, d2, info2 = ref.step(a)
        np.testing.assert_array_equal(o, o2, err_msg=f"{case}: observation at step {t}")
        np.testing.assert_array_equal(r, r2, err_msg=f"{case}: weighted reward at step {t}")
        np.testing.assert_array_equal(d, d2, err_msg=f"{case}: done at step {t}")
        n_done += int(d.sum())
    assert n_done > 2 * len(rotation), "the rotation list was meant to wrap"
    env.close()


def test_cycle_maps_with_both_env_kinds_are_refused():
    with pytest.raises((NotImplementedError, ValueError)):
        make_env("hostsim", 2, 1, ai2s=[P], map_paths=[MAP_8], cycle_maps=[MAP_8])


# Long scripted-bot games (many path searches, full-length episodes, several resets), hostsim only: the
# CUDA build runs the same cases through the shorter table above and test_gpu_api's large slabs.
STRESS = {
    # name: (map, num_selfplay, ai2s, partial_obs, max_steps, steps, eps, seed)
    "wrush16_long": (MAP_16, 0, [W] * 12, False, 2000, 2300, 0.02, 1),
    "wrush8_long": (MAP_8, 0, [W] * 16, False, 600, 1300, 0.05, 2),
    "wrush16_partial_long": (MAP_16, 0, [W] * 8, True, 1000, 1200, 0.02, 3),
    "mix10_long": ("maps/10x10/basesWorkers10x10.xml", 4, [W, LR] * 5, False, 800, 1200, 0.05, 4),
    "lrush16_long": (MAP_16, 0, [LR] * 8, False, 2000, 2300, 0.02, 5),
}


@pytest.mark.parametrize("case", list(STRESS))
def test_long_scripted_bot_games_match_oracle(case):
    mp, nsp, ais, partial, max_steps, steps, eps, seed = STRESS[case]
    kw = dict(num_selfplay_envs=nsp, num_bot_envs=len(ais), partial_obs=partial, max_steps=max_steps, ai2s=ais, map_paths=[mp], reward_weight=RW)
    ref = make_env("oracle", **kw)
    env = make_env("hostsim", **kw)
    rng = np.random.default_rng(seed)
    np.testing.assert_array_equal(env.reset(), ref.reset())
    for t in range(steps):
        m_ref = ref.get_action_mask()
        np.testing.assert_array_equal(env.get_action_mask(), m_ref, err_msg=f"{case}: mask at step {t}")
        a = sample_actions(m_ref, rng, eps)
        o, r, d, _ = env.step(a)
        o2, r2, d2, _ = ref.step(a)
        np.testing.assert_array_equal(o, o2, err_msg=f"{case}: observation at step {t}")
        np.testing.assert_array_equal(r, r2, err_msg=f"{case}: reward at step {t}")
        np.testing.assert_array_equal(d, d2, err_msg=f"{case}: done at step {t}")


# Bot vs bot (J/tests/JNIBotClient.java:108-134 under MicroRTSBotVecEnv): both players
# scripted, both acting on the full state, player 0's action issued first.  The reference returns only
# rewards / dones; the observation of the same state is compared as well (it exposes every unit).
BOTVBOT = {
    # name: (map, ai1s, ai2s, max_steps, steps)
    "wr_vs_wr16": (MAP_16, [W, W], [W, W], 600, 700),
    "mixed8": (MAP_8, [W, LR, P, W, LR, LR], [LR, W, W, P, LR, P], 300, 700),
    "lr_vs_wr16": (MAP_16, [LR, W, LR], [W, LR, LR], 1500, 1700),
    "heavy_ranged16": (MAP_16, [POH, POR, POL, W], [POR, POH, POR, POH], 1500, 1700),
    "nobase8": ("maps/8x8/noBases8x8.xml", [LR, W], [W, LR], 250, 400),
}


@pytest.mark.parametrize("engine", _engines())
@pytest.mark.parametrize("case", list(BOTVBOT))
def test_bot_vs_bot_games_match_oracle(engine, case):
    mp, ai1s, ai2s, max_steps, steps = BOTVBOT[case]
    kw = dict(num_selfplay_envs=0, num_bot_envs=len(ai2s), max_steps=max_steps, ai2s=ai2s, ai1s=ai1s, map_paths=[mp], reward_weight=RW)
    ref = make_env("oracle", **kw)
    env = make_env(engine, **kw)
    np.testing.assert_array_equal(env.reset(), ref.reset())
    n_done = 0
    a = np.zeros((len(ai2s), ref.height * ref.width, 7), np.int32)  # ignored by both
    for t in range(steps):
        o, r, d, info = env.step(a)
        o2, r2, d2, info2 = ref.step(a)
        np.testing.assert_array_equal(np.stack([i["raw_rewards"] for i in info]), np.stack([i["raw_rewards"] for i in info2]), err_msg=f"{case}: raw rewards at step {t}")
        np.testing.assert_array_equal(d, d2, err_msg=f"{case}: done at step {t}")
        np.testing.assert_array_equal(o, o2, err_msg=f"{case}: observation at step {t}")
        n_done += int(d.sum())
    assert n_done > 0
    env.close()
