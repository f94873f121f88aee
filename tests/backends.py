"""TEST INFRASTRUCTURE.  Environment factories with the reference's class surface
(gym_microrts/envs/vec_env.py: reset / get_action_mask / step) over three engines:

  "oracle"   the CPU restatement of the Java path (oracle/)                 -- the checker
  "hostsim"  the device rule/encode headers compiled for the host (CPU)     -- logic debugging
  "cuda"     microrts_py_b200.envs.MicroRTSGridModeVecEnv (libmrts_b200.so)  -- the product

so that the known-answer tests read like the reference's own tests/test_*.py.
"""
import itertools

import numpy as np

from microrts_py_b200 import microrts_ai
from tests.common import templates_for

AI_IDS = {microrts_ai.passiveAI: 0, microrts_ai.workerRushAI: 1, microrts_ai.lightRushAI: 2, microrts_ai.POWorkerRush: 3, microrts_ai.POLightRush: 4,
          microrts_ai.POHeavyRush: 5, microrts_ai.PORangedRush: 6}


class _HostEnv:
    """Reference-shaped wrapper (vec_env.py:278-282, 1001-1061, 1091-1101) over a raw engine."""

    def __init__(self, engine, num_envs, h, w, reward_weight, reward_shaping=True, num_bot_envs=0, cycle_templates=()):
        self.e = engine
        # oracle only: the rotation loop of the reference's step_wait runs here, over the engine's clients
        self.num_bot_envs = num_bot_envs
        self.next_map = itertools.cycle(cycle_templates) if len(cycle_templates) else None
        self.num_envs, self.height, self.width = num_envs, h, w
        self.reward_weight = np.asarray(reward_weight, np.float64)
        self.reward_shaping = reward_shaping
        self.source_unit_mask = None

    def reset(self):
        return self.e.reset()

    def get_action_mask(self):
        m = self.e.get_masks()[0]
        self.source_unit_mask = m[:, :, 0]
        return m

    def step(self, actions):
        obs, reward, done = self.e.step(np.asarray(actions).reshape(self.num_envs, self.height * self.width, 7))[:3]
        if not self.reward_shaping:
            reward[:, 1:] = 0
        if self.next_map is not None:
            self.e.cycle_maps_after_step(obs, done, self.num_bot_envs, self.next_map)
        infos = [{"raw_rewards": item} for item in reward]
        return obs, reward @ self.reward_weight, done.astype(bool), infos

    def close(self):
        pass


def make_env(backend, num_selfplay_envs=0, num_bot_envs=0, partial_obs=False, max_steps=2000, ai2s=(), map_paths=(), reward_weight=(0.0, 1.0, 0.0, 0.0, 0.0, 5.0), cycle_maps=(), ai1s=None, **kw):
    n = num_selfplay_envs + num_bot_envs
    paths = list(map_paths) * n if len(map_paths) == 1 else list(map_paths)
    if backend == "cuda":
        from microrts_py_b200.envs import MicroRTSGridModeVecEnv

        return MicroRTSGridModeVecEnv(
            num_selfplay_envs, num_bot_envs, partial_obs=partial_obs, max_steps=max_steps, ai2s=list(ai2s), map_paths=paths, reward_weight=np.asarray(reward_weight), cycle_maps=list(cycle_maps), ai1s=ai1s, **kw
        )
    t = templates_for(paths)
    ids = [AI_IDS[a] for a in ai2s]
    ids1 = [AI_IDS[a] for a in ai1s] if ai1s is not None else None  # bot vs bot (JNIBotClient)
    if backend == "oracle":
        from oracle.oracle_py import OracleVecEnv

        eng = OracleVecEnv(num_selfplay_envs, num_bot_envs, max_steps, partial_obs, t, ids, ai1_ids=ids1)
        return _HostEnv(eng, n, t[0].height, t[0].width, reward_weight, num_bot_envs=num_bot_envs, cycle_templates=templates_for(list(cycle_maps)))
    elif backend == "hostsim":
        from tests.hostsim import HostSimVecEnv

        eng = HostSimVecEnv(num_selfplay_envs, num_bot_envs, max_steps, partial_obs, t, ids, cycle_templates=templates_for(list(cycle_maps)), ai1_ids=ids1)
    else:
        raise ValueError(backend)
    return _HostEnv(eng, n, t[0].height, t[0].width, reward_weight)
