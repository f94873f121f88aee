"""Parity of every kernel instantiation and of the BASELINE configs against the oracle (round-2 matrix).

Each case names the product configuration it exercises on the device:
  pipeline  "direct" = rules pass writes the tensors (mrts_rules_warp_kernel / mrts_rules_block_kernel, plain stores)
            "split"  = mrts_rules_warp_kernel -> packed words / entries -> mrts_expand_kernel<ELEM, PLANES> (bulk stores)
  dtype     int32 / uint8 (ELEM 4 / 1), partial_obs (PLANES 31 / 29), map size (warp kernels up to 32x32 in direct
            mode, the 256-thread block kernel on 64x64), scripted players (light / heavy kernels, bot vs bot).
"hostsim" runs the same rule / encode headers on the CPU (direct form; MRTS_HOSTSIM_SPLIT=1 for the packed form).
The reference semantics under test: JNIGridnetClient(SelfPlay).gameStep, JNIBotClient.gameStep (:108-134),
PartiallyObservableGameState (:23-58), JNIGridnetVecClient.gameStep terminal handling (:213-297)."""
import numpy as np
import pytest

from microrts_py_b200 import microrts_ai
from tests.backends import make_env
from tests.common import MAP_4, MAP_8, MAP_16, MAP_32, MAP_64, sample_actions

P, W, LR = microrts_ai.passiveAI, microrts_ai.workerRushAI, microrts_ai.lightRushAI
POW, POL, POH, POR = microrts_ai.POWorkerRush, microrts_ai.POLightRush, microrts_ai.POHeavyRush, microrts_ai.PORangedRush
RW = np.array([10.0, 1.0, 1.0, 0.2, 1.0, 4.0])
MAP_ODD = "maps/itsNotSafe.xml"       # 15 x 14 = 210 cells: env blocks of the int32 tensors are not 16-byte aligned (scalar store tail)
MAP_9x8 = "maps/NoWhereToRun9x8.xml"  # 72 cells: aligned for int32, not for uint8
MAP_10 = "maps/10x10/basesWorkers10x10.xml"


def run_case(engine, nsp, ais, mp, partial, max_steps, steps, eps, seed=1234, ai1s=None, every=1, cuda_kw=None, expect_done=True, expect_deaths=False):
    """Same seeded action trace into the oracle and the engine; obs / mask compared every `every` steps (and on every
    step that ends an episode), raw rewards and dones on every step."""
    kw = dict(num_selfplay_envs=nsp, num_bot_envs=len(ais), partial_obs=partial, max_steps=max_steps, ai2s=ais, map_paths=[mp], reward_weight=RW, ai1s=ai1s)
    ref = make_env("oracle", **kw)
    env = make_env(engine, **kw, **(cuda_kw or {})) if engine == "cuda" else make_env(engine, **kw)
    rng = np.random.default_rng(seed)
    n = nsp + len(ais)
    agents = ai1s is None or any(a is None for a in ai1s)

    def i32(x):
        return np.asarray(x).astype(np.int32)

    np.testing.assert_array_equal(i32(env.reset()), ref.reset(), err_msg="reset observation")
    n_done, min_units, prev_units = 0, 10 ** 9, None
    deaths = 0
    zero_actions = np.zeros((n, ref.height * ref.width, 7), np.int32)
    for t in range(steps):
        if agents:
            m_ref = ref.get_action_mask()
            if t % every == 0:
                np.testing.assert_array_equal(i32(env.get_action_mask()), m_ref, err_msg=f"mask at step {t}")
            a = sample_actions(m_ref, rng, eps)
        else:
            a = zero_actions
        o, r, d, info = env.step(a)
        o2, r2, d2, info2 = ref.step(a)
        np.testing.assert_array_equal(np.stack([i["raw_rewards"] for i in info]), np.stack([i["raw_rewards"] for i in info2]), err_msg=f"raw rewards at step {t}")
        np.testing.assert_array_equal(d, d2, err_msg=f"done at step {t}")
        if t % every == 0 or d2.any():
            np.testing.assert_array_equal(i32(o), o2, err_msg=f"observation at step {t}")
        n_done += int(d2.sum())
        units = (o2[..., 13] == 0).sum(axis=(1, 2))  # cells whose unit-type plane 0 ("no unit") is off
        if prev_units is not None:
            deaths += int(np.maximum(prev_units - units, 0)[~d2].sum())
        prev_units = units
    if expect_done:
        assert n_done > 0, "the case was meant to cross an episode end"
    if expect_deaths:
        assert deaths > 0, "the case was meant to see units die"
    env.close()
    return n_done


# ---------------------------------------------------------------------------------------------------------------
# (a) 64x64 that fights: bot vs bot on BloodBath, two full games per env decided by elimination (~2,400 frames each)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["direct", "block", "split"])
def test_bloodbath64_bot_games_fight_to_the_end(pipeline):
    ckw = dict(pipeline="direct", threads_per_game=256) if pipeline == "block" else dict(pipeline=pipeline, chunks=2)
    n = run_case("cuda", 0, [W, LR], MAP_64, False, 6000, 5200, 0.0, ai1s=[LR, W], every=16, cuda_kw=ckw, expect_deaths=True)
    assert n >= 2


def test_bloodbath64_bot_games_hostsim_prefix():
    run_case("hostsim", 0, [W, LR], MAP_64, False, 6000, 1400, 0.0, ai1s=[LR, W], every=8, expect_done=False, expect_deaths=True)


# ---------------------------------------------------------------------------------------------------------------
# (b) element size x plane count x map size x pipeline
# ---------------------------------------------------------------------------------------------------------------
MATRIX = {
    # name: (nsp, ais, map, partial, max_steps, steps, eps, cuda kwargs)
    "u8_p31_8x8_direct": (4, [W, LR, P], MAP_8, True, 150, 500, 0.08, dict(dtype=np.uint8, pipeline="direct")),
    "u8_p31_8x8_split": (4, [W, LR, P], MAP_8, True, 150, 500, 0.08, dict(dtype=np.uint8, pipeline="split", chunks=2)),
    "u8_p31_32x32_direct": (2, [W, POL], MAP_32, True, 200, 450, 0.05, dict(dtype=np.uint8, pipeline="direct")),
    "u8_p31_32x32_warp": (2, [W, POL], MAP_32, True, 200, 450, 0.05, dict(dtype=np.uint8, pipeline="direct", threads_per_game=-2)),
    "i32_p29_32x32_block128": (2, [LR, P], MAP_32, False, 200, 400, 0.05, dict(pipeline="direct")),
    "u8_p31_32x32_split": (2, [W, POL], MAP_32, True, 200, 450, 0.05, dict(dtype=np.uint8, pipeline="split", chunks=3)),
    "u8_p31_64x64_block": (2, [W], MAP_64, True, 120, 300, 0.05, dict(dtype=np.uint8, pipeline="direct", threads_per_game=256)),
    "u8_p31_64x64_warp": (2, [W], MAP_64, True, 120, 300, 0.05, dict(dtype=np.uint8, pipeline="direct")),
    "u8_p31_64x64_split": (2, [W], MAP_64, True, 120, 300, 0.05, dict(dtype=np.uint8, pipeline="split", chunks=2)),
    "u8_p29_64x64_block": (2, [LR], MAP_64, False, 150, 320, 0.05, dict(dtype=np.uint8, pipeline="direct", threads_per_game=256)),
    "i32_p31_64x64_block": (2, [POW], MAP_64, True, 150, 320, 0.05, dict(pipeline="direct", threads_per_game=256)),
    "i32_p31_64x64_warp": (2, [POW], MAP_64, True, 150, 320, 0.05, dict(pipeline="direct")),
    "i32_p29_16x16_plain_stores": (4, [W, P], MAP_16, False, 200, 500, 0.05, dict(pipeline="direct", threads_per_game=64)),
    "i32_p29_64x64_split": (2, [W], MAP_64, False, 150, 320, 0.05, dict(pipeline="split", chunks=2)),
    "i32_p29_16x16_split4": (6, [P, W, LR, W], MAP_16, False, 250, 700, 0.05, dict(pipeline="split", chunks=4)),
    "i32_p31_16x16_split": (4, [POW, POH, P], MAP_16, True, 300, 650, 0.05, dict(pipeline="split", chunks=2)),
    "u8_p29_16x16_split": (4, [W, P], MAP_16, False, 200, 500, 0.05, dict(dtype=np.uint8, pipeline="split", chunks=2)),
    # 64-register light kernel and 96-register heavy kernel side by side on the smallest map
    "i32_p29_4x4_split": (2, [P, W], MAP_4, False, 50, 300, 0.2, dict(pipeline="split", chunks=2)),
    # (c) odd areas: 210 cells -> every second env block of the int32 tensors is unaligned (scalar stores); 72 / 100 cells
    # are aligned for int32 only: uint8 falls back to the direct pass even when the split pipeline is asked for
    "i32_odd15x14_direct": (2, [W, LR, P], MAP_ODD, False, 200, 500, 0.05, dict(pipeline="direct")),
    "i32_odd15x14_split_falls_back": (2, [W, LR, P], MAP_ODD, False, 200, 500, 0.05, dict(pipeline="split")),
    "u8_odd15x14_partial": (2, [W, P], MAP_ODD, True, 200, 400, 0.05, dict(dtype=np.uint8)),
    "i32_9x8_split": (2, [W, LR], MAP_9x8, False, 150, 400, 0.05, dict(pipeline="split", chunks=2)),
    # (d) host-buffer transport with result-buffer slots: device-side delta compaction (mrts_delta_kernel) + host patching;
    # 5 x 210 cells is not a multiple of 32 (partial bitmap word), resets rewrite whole maps
    "host_i32_odd15x14": (2, [W, LR, P], MAP_ODD, False, 200, 500, 0.05, dict(host_transport=True)),
    "host_u8_p31_8x8": (4, [W, LR, P], MAP_8, True, 150, 500, 0.08, dict(dtype=np.uint8, host_transport=True)),
    "host_i32_p31_32x32": (2, [W, POL], MAP_32, True, 200, 300, 0.05, dict(host_transport=True)),
    "u8_9x8_split_falls_back": (2, [W, LR], MAP_9x8, True, 150, 400, 0.05, dict(dtype=np.uint8, pipeline="split")),
    "u8_10x10": (4, [W, LR], MAP_10, False, 200, 450, 0.05, dict(dtype=np.uint8)),
}


@pytest.mark.gpu
@pytest.mark.parametrize("case", list(MATRIX))
def test_instantiation_matrix_matches_oracle(case):
    nsp, ais, mp, partial, max_steps, steps, eps, ckw = MATRIX[case]
    run_case("cuda", nsp, ais, mp, partial, max_steps, steps, eps, cuda_kw=ckw)


@pytest.mark.parametrize("case", ["i32_odd15x14_direct", "i32_9x8_split", "u8_p31_32x32_direct"])
def test_matrix_maps_on_hostsim(case):
    nsp, ais, mp, partial, max_steps, steps, eps, _ = MATRIX[case]
    run_case("hostsim", nsp, ais, mp, partial, max_steps, min(steps, 300), eps, expect_done=max_steps < 300)


# ---------------------------------------------------------------------------------------------------------------
# (d) the long scripted-bot games (full-length episodes, several resets) on the device as well
# ---------------------------------------------------------------------------------------------------------------
from tests.test_trajectory_parity import STRESS  # noqa: E402


@pytest.mark.gpu
@pytest.mark.parametrize("case", list(STRESS))
def test_long_scripted_bot_games_match_oracle_on_the_device(case):
    mp, nsp, ais, partial, max_steps, steps, eps, seed = STRESS[case]
    run_case("cuda", nsp, ais, mp, partial, max_steps, steps, eps, seed=seed, every=4)


# ---------------------------------------------------------------------------------------------------------------
# (e) the literal BASELINE config 1: N = 2 self-play views, max_steps = 2000, 4,000 steps (two truncations)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_config1_literally():
    assert run_case("cuda", 2, [], MAP_16, False, 2000, 4000, 0.03, every=2) >= 2


@pytest.mark.gpu
def test_config2_literally():
    """ppo_gridnet rollout shape: 24 bot envs, 12 vs passiveAI and 12 vs workerRushAI, 2,048 steps."""
    run_case("cuda", 0, [P] * 12 + [W] * 12, MAP_16, False, 2000, 2048, 0.03, every=4)


# ---------------------------------------------------------------------------------------------------------------
# explicit reset() in the middle of an episode: envSteps is NOT cleared (JNIGridnetVecClient.reset, :179-211), so the
# first episode after it is truncated early
# ---------------------------------------------------------------------------------------------------------------
def _mid_episode_reset(engine):
    kw = dict(num_selfplay_envs=2, num_bot_envs=2, max_steps=60, ai2s=[P, W], map_paths=[MAP_8], reward_weight=RW)
    ref, env = make_env("oracle", **kw), make_env(engine, **kw)
    rng = np.random.default_rng(5)
    np.testing.assert_array_equal(env.reset(), ref.reset())
    dones = []
    for t in range(150):
        if t == 25:
            np.testing.assert_array_equal(env.reset(), ref.reset(), err_msg="mid-episode reset")
        m = ref.get_action_mask()
        np.testing.assert_array_equal(env.get_action_mask(), m)
        a = sample_actions(m, rng, 0.05)
        o, r, d, _ = env.step(a)
        o2, r2, d2, _ = ref.step(a)
        np.testing.assert_array_equal(o, o2, err_msg=f"step {t}")
        np.testing.assert_array_equal(d, d2, err_msg=f"done at step {t}")
        if d2.all():
            dones.append(t)
    assert dones[0] == 59, "the counter kept running through the explicit reset: first truncation at frame 60, not 25 + 60"
    env.close()


def test_explicit_reset_keeps_the_step_counter_hostsim():
    _mid_episode_reset("hostsim")


@pytest.mark.gpu
def test_explicit_reset_keeps_the_step_counter_cuda():
    _mid_episode_reset("cuda")
