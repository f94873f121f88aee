"""The reference's own known-answer tests (tests/test_observation.py, tests/test_mask.py,
tests/test_reward.py of adFrej/MicroRTS-Py), run against the oracle (pins it), the host build of
the device headers, and -- on the GPU box -- the CUDA path through MicroRTSGridModeVecEnv.
Vectors come from tests/golden/known_answers.json (made by tests/golden/make_known_answers.py
from the reference's test files)."""
import numpy as np

from microrts_py_b200 import microrts_ai
from tests.backends import make_env

RW = np.array([10.0, 1.0, 1.0, 0.2, 1.0, 4.0])


def test_observation(backend, known_answers):  # reference tests/test_observation.py:8-108
    ka = known_answers["obs"]
    cells = {k: np.asarray(v, np.int32) for k, v in ka["cells"].items()}
    by_map = {}
    for m, env, y, x, name in ka["checks"]:
        by_map.setdefault(m, []).append((env, y, x, name))
    for m, checks in by_map.items():
        envs = make_env(backend, num_selfplay_envs=2, max_steps=5000, map_paths=[m], reward_weight=RW)
        next_obs = envs.reset()
        assert next_obs.dtype == np.int32 and next_obs.shape[0] == 2 and next_obs.shape[3] == 29
        for env, y, x, name in checks:
            np.testing.assert_array_equal(next_obs[env][y][x], cells[name], err_msg=f"{m} env {env} ({y},{x}) {name}")
        if "16x16" in m:
            assert next_obs.sum() == ka["total"] == 3072
        assert (next_obs.sum(axis=3) == 6).all()  # exactly one plane per feature group in every cell
        envs.close()


def _mask_env(backend):
    return make_env(backend, num_bot_envs=1, max_steps=2000, ai2s=[microrts_ai.passiveAI], map_paths=["maps/4x4/baseTwoWorkers4x4.xml"], reward_weight=RW)


def test_mask(backend, known_answers):  # reference tests/test_mask.py:9-85
    envs = _mask_env(backend)
    envs.reset()
    mask = np.array(envs.get_action_mask())
    assert mask.shape == (1, 16, 78) and mask.dtype == np.int32
    for cell, vec in known_answers["mask"]["rows"]:
        np.testing.assert_array_equal(mask[0, cell], np.asarray(vec, np.int32), err_msg=f"cell {cell}")
    others = [c for c in range(16) if c not in (1, 4, 5)]
    assert mask[0, others].sum() == 0
    envs.close()


def _act(envs, cell, vec):
    action = np.zeros(envs.height * envs.width * 7, np.int32)
    action[cell * 7 : (cell + 1) * 7] = vec
    return action


def test_reward(backend):  # reference tests/test_reward.py:9-48
    envs = _mask_env(backend)
    envs.reset()
    envs.get_action_mask()
    assert envs.step(_act(envs, 1, [2, 0, 3, 0, 0, 0, 0]))[1].flatten() > 0  # mine
    for _ in range(20):  # wait for the action to finish
        envs.get_action_mask()
        envs.step(np.zeros(16 * 7, np.int32))
    envs.get_action_mask()
    assert envs.step(_act(envs, 1, [3, 0, 0, 2, 0, 0, 0]))[1].flatten() > 0  # return
    envs.close()


def test_produce_worker_reward(backend):  # reference tests/test_reward.py:51-76
    envs = _mask_env(backend)
    envs.reset()
    envs.get_action_mask()
    obs, r, d, infos = envs.step(_act(envs, 5, [4, 0, 0, 0, 1, 3, 0]))
    assert r.flatten() > 0
    np.testing.assert_array_equal(infos[0]["raw_rewards"], [0, 0, 1, 0, 0, 0])
    envs.close()


def test_produce_buildings_reward(backend):  # reference tests/test_reward.py:79-104
    envs = _mask_env(backend)
    envs.reset()
    envs.get_action_mask()
    obs, r, d, infos = envs.step(_act(envs, 4, [4, 0, 0, 0, 2, 2, 0]))
    assert r.flatten() > 0
    np.testing.assert_array_equal(infos[0]["raw_rewards"], [0, 0, 0, 1, 0, 0])
    envs.close()
