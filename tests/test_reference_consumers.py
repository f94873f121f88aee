"""The drop-in under the reference's own consumers.

`StatsRecorder` restates MicroRTSStatsRecorder.step_wait (experiments/ppo_gridnet.py:126-158) operation for operation --
`list(infos[:])`, `infos[i]["raw_rewards"]`, `.sum()`, `infos[i].copy()`, `str(rf) for rf in self.rfs`, `newinfos[i] = info`
-- and `rollout` is the env-facing part of the PPO loop (:459-480: get_action_mask -> policy -> `step(action.reshape(N, -1))`,
`for info in infos: "episode" in info.keys()`).  Every mode of the env class has to survive them and produce the same episode
statistics as the oracle driven by the same code.  The CPU half checks the list-like `infos` object itself."""
import numpy as np
import pytest

from microrts_py_b200 import microrts_ai
from tests.backends import make_env
from tests.common import MAP_16, sample_actions

P, W = microrts_ai.passiveAI, microrts_ai.workerRushAI


class StatsRecorder:
    def __init__(self, env, gamma=0.99):
        self.venv, self.gamma = env, gamma
        self.num_envs = env.num_envs
        self.rfs = getattr(env, "rfs", None) or ["WinLossRewardFunction", "ResourceGatherRewardFunction", "ProduceWorkerRewardFunction",
                                                   "ProduceBuildingRewardFunction", "AttackRewardFunction", "ProduceCombatUnitRewardFunction"]

    def reset(self):
        obs = self.venv.reset()
        self.raw_rewards = [[] for _ in range(self.num_envs)]
        self.ts = np.zeros(self.num_envs, dtype=np.float32)
        self.raw_discount_rewards = [[] for _ in range(self.num_envs)]
        return obs

    def get_action_mask(self):
        return self.venv.get_action_mask()

    def step(self, actions):
        obs, rews, dones, infos = self.venv.step(actions)
        newinfos = list(infos[:])
        for i in range(len(dones)):
            self.raw_rewards[i] += [infos[i]["raw_rewards"]]
            self.raw_discount_rewards[i] += [(self.gamma ** self.ts[i]) * np.concatenate((infos[i]["raw_rewards"], infos[i]["raw_rewards"].sum()), axis=None)]
            self.ts[i] += 1
            if dones[i]:
                info = infos[i].copy()
                raw_returns = np.array(self.raw_rewards[i]).sum(0)
                raw_names = [str(rf) for rf in self.rfs]
                raw_discount_returns = np.array(self.raw_discount_rewards[i]).sum(0)
                raw_discount_names = ["discounted_" + str(rf) for rf in self.rfs] + ["discounted"]
                info["microrts_stats"] = dict(zip(raw_names, raw_returns))
                info["microrts_stats"].update(dict(zip(raw_discount_names, raw_discount_returns)))
                self.raw_rewards[i] = []
                self.raw_discount_rewards[i] = []
                self.ts[i] = 0
                newinfos[i] = info
        return obs, rews, dones, newinfos


def rollout(envs, ref, steps, seed):
    """ppo_gridnet.py:459-480 with the masked-uniform stand-in policy; returns the episode statistics seen."""
    import torch

    def host(x):
        return x.cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)

    rng = np.random.default_rng(seed)
    next_obs, ref_obs = host(envs.reset()), ref.reset()
    np.testing.assert_array_equal(next_obs.astype(np.int32), ref_obs)
    stats, ref_stats = [], []
    for step in range(steps):
        mask = host(envs.get_action_mask())
        ref_mask = ref.get_action_mask()
        np.testing.assert_array_equal(mask.astype(np.int32), ref_mask, err_msg=f"mask at step {step}")
        action = sample_actions(ref_mask, rng, 0.03)
        next_obs, rs, ds, infos = envs.step(action.reshape(envs.num_envs, -1))
        ref_obs, ref_rs, ref_ds, ref_infos = ref.step(action.reshape(envs.num_envs, -1))
        np.testing.assert_array_equal(host(next_obs).astype(np.int32), ref_obs, err_msg=f"obs at step {step}")
        np.testing.assert_allclose(host(rs), ref_rs, rtol=0, atol=1e-12)
        np.testing.assert_array_equal(host(ds), ref_ds)
        assert isinstance(infos, list) and len(infos) == envs.num_envs
        for info, rinfo in zip(infos, ref_infos):
            assert ("microrts_stats" in info.keys()) == ("microrts_stats" in rinfo.keys())
            if "microrts_stats" in info.keys():
                stats.append(info["microrts_stats"])
                ref_stats.append(rinfo["microrts_stats"])
    assert len(stats) > 0, "the rollout was meant to finish episodes"
    for a, b in zip(stats, ref_stats):
        assert a.keys() == b.keys()
        for k in a:
            assert a[k] == pytest.approx(b[k], abs=1e-9), k
    return stats


MODES = {
    "numpy": dict(),
    "torch": dict(return_torch=True),
    "host_transport": dict(host_transport=True),
    "zero_copy": dict(zero_copy=True),
    "uint8_split": dict(dtype=np.uint8, pipeline="split", chunks=2),
}


@pytest.mark.gpu
@pytest.mark.parametrize("mode", list(MODES))
def test_stats_recorder_and_rollout_loop_over_every_mode(mode):
    kw = dict(num_selfplay_envs=2, num_bot_envs=4, max_steps=120, ai2s=[P, W, W, P], map_paths=[MAP_16], reward_weight=np.array([10.0, 1.0, 1.0, 0.2, 1.0, 4.0]))
    envs = StatsRecorder(make_env("cuda", **kw, **MODES[mode]))
    ref = StatsRecorder(make_env("oracle", **kw))
    stats = rollout(envs, ref, 400, seed=11)
    assert "WinLossRewardFunction" in stats[0] and "discounted" in stats[0]
    envs.venv.close()


def test_stats_recorder_over_hostsim():
    """the consumer restatement itself, without a GPU"""
    kw = dict(num_selfplay_envs=2, num_bot_envs=2, max_steps=80, ai2s=[P, W], map_paths=[MAP_16], reward_weight=np.array([10.0, 1.0, 1.0, 0.2, 1.0, 4.0]))
    rollout(StatsRecorder(make_env("hostsim", **kw)), StatsRecorder(make_env("oracle", **kw)), 200, seed=3)


def test_infos_object_behaves_like_the_reference_list():
    """RawRewardInfos (torch / host_transport modes) under the operations the reference's wrappers apply to `infos`."""
    from microrts_py_b200.envs.vec_env import RawRewardInfos

    raw = np.arange(24, dtype=np.float64).reshape(4, 6)
    infos = RawRewardInfos(raw.copy())
    assert len(infos) == 4
    as_list = list(infos[:])
    assert isinstance(as_list, list) and len(as_list) == 4 and all(isinstance(i, dict) for i in as_list)
    np.testing.assert_array_equal(as_list[2]["raw_rewards"], raw[2])
    as_list[1] = {"replaced": True}  # `newinfos[i] = info`
    assert "raw_rewards" in infos[1]
    c = infos[3].copy()
    c["episode"] = 1
    assert "episode" not in infos[3]
    assert infos[0]["raw_rewards"].sum() == raw[0].sum()
    assert [i["raw_rewards"][0] for i in infos] == [0.0, 6.0, 12.0, 18.0]
    assert [i["raw_rewards"][0] for i in infos[1:3]] == [6.0, 12.0]
    assert infos[-1]["raw_rewards"][5] == 23.0
