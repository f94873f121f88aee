"""GPU-only checks of the product surface beyond the reference-shaped NumPy path: torch
(device-resident) mode, uint8 element mode, the device action sampler, un-fused masks, golden
trace replay through the raw engine entry points, and full-size property checks."""
import ctypes as C

import numpy as np
import pytest

from microrts_py_b200 import microrts_ai
from tests.common import GROUPS, MAP_8, MAP_16, MAP_32, sample_actions, templates_for

pytestmark = pytest.mark.gpu
P, W, LR = microrts_ai.passiveAI, microrts_ai.workerRushAI, microrts_ai.lightRushAI


def _env(*a, **k):
    from microrts_py_b200.envs import MicroRTSGridModeVecEnv

    return MicroRTSGridModeVecEnv(*a, **k)


def _oracle(nsp, ais, mp, partial=False, max_steps=2000):
    from oracle.oracle_py import OracleVecEnv

    n = nsp + len(ais)
    return OracleVecEnv(nsp, len(ais), max_steps, partial, templates_for([mp] * n), [a(None).ai_id for a in ais])


@pytest.mark.parametrize("mode", ["torch", "uint8", "unfused", "zero_copy", "host_transport", "host_transport_uint8"])
def test_modes_match_oracle(mode):
    import torch

    kw = dict(torch=dict(return_torch=True), uint8=dict(dtype=np.uint8), unfused=dict(fuse_mask=False), zero_copy=dict(zero_copy=True),
              host_transport=dict(host_transport=True), host_transport_uint8=dict(host_transport=True, dtype=np.uint8))[mode]
    ais = [P, W]
    env = _env(4, 2, max_steps=150, ai2s=ais, map_paths=[MAP_16], **kw)
    ref = _oracle(4, ais, MAP_16, max_steps=150)
    rng = np.random.default_rng(7)

    def host(x):
        return x.cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)

    np.testing.assert_array_equal(host(env.reset()).astype(np.int32), ref.reset())
    for t in range(320):
        m = host(env.get_action_mask())
        m_ref, src_ref = ref.get_masks()
        np.testing.assert_array_equal(m.astype(np.int32), m_ref, err_msg=f"mask {t}")
        np.testing.assert_array_equal(host(env.source_unit_mask).astype(np.int32), src_ref)
        a = sample_actions(m_ref, rng, 0.05)
        act = torch.as_tensor(a.astype(np.int64)).cuda() if mode == "torch" else a.astype(np.int64)
        o, r, d, infos = env.step(act)
        o2, r2, d2 = ref.step(a)
        np.testing.assert_array_equal(host(o).astype(np.int32), o2, err_msg=f"obs {t}")
        np.testing.assert_array_equal(host(d).astype(np.uint8), d2)
        np.testing.assert_array_equal(host(infos[3]["raw_rewards"]), r2[3])
        np.testing.assert_allclose(host(r), r2 @ env.reward_weight, rtol=0, atol=1e-12)
    env.close()


def test_device_sampler_draws_from_the_mask_and_is_reproducible():
    import torch

    runs = []
    for rep in range(2):
        env = _env(64, 8, max_steps=100, ai2s=[P, W] * 4, map_paths=[MAP_8], return_torch=True, seed=5)
        env.reset()
        hist = []
        for t in range(150):
            m = env.get_action_mask()
            a = env.sample_actions()
            src = m[:, :, 0].bool()
            assert src.any()
            av, mv = a[src].long(), m[src]
            for k, (lo, hi) in enumerate(GROUPS):
                grp = mv[:, lo:hi]
                chosen = torch.gather(grp, 1, av[:, k : k + 1]).squeeze(1)
                empty = grp.sum(1) == 0
                assert bool(((chosen == 1) | (empty & (av[:, k] == 0))).all()), f"component {k} step {t}"
            o, r, d, _ = env.step(a)
            hist.append((o.sum().item(), r.sum().item(), d.sum().item()))
        runs.append(hist)
        env.close()
    assert runs[0] == runs[1]
    assert any(h[2] > 0 for h in runs[0])


def test_golden_trace_replay_on_device(golden_traces):
    """TestTracesIntegrity.java:72-126 strengthened to state equality, through mrts_debug_issue /
    mrts_debug_cycle on the GPU: a subset of the reference's 280 traces incl. the config maps."""
    from microrts_py_b200 import _capi
    from microrts_py_b200.maps import load_map

    L = _capi.lib()
    want = ("basesWorkers16x16A", "basesWorkers8x8", "BWDistantResources32x32", "BloodBath", "melee", "4x4")
    picked = [t for t in golden_traces if any(w in t["map"] for w in want)]
    assert len(picked) >= 8
    for tr in picked[:24]:
        tmpl = load_map(tr["map"])
        cfg, keep = _capi.make_config(2, 0, 100000, False, [tmpl, tmpl], [], none_mode=_capi.NONE_UPSTREAM)
        h = C.c_void_p()
        _capi.check(L.mrts_create(C.byref(cfg), C.byref(h)))
        cap = L.mrts_unit_capacity(h)
        hdr, rows = np.zeros(5, np.int32), np.zeros((cap, 12), np.int32)
        time = 0
        for k, e in enumerate(tr["entries"]):
            while time < e[0]:
                _capi.check(L.mrts_debug_cycle(h, 0, None, None))
                time += 1
            n = _capi.check(L.mrts_export_state(h, 0, hdr.ctypes.data_as(C.c_void_p), rows.ctypes.data_as(C.c_void_p), cap, None))
            units = np.asarray(e[3], np.int32).reshape(-1, 6)
            assert (hdr[0], hdr[1], hdr[2]) == (e[0], e[1], e[2]), f"{tr['map']} entry {k}"
            np.testing.assert_array_equal(rows[:n, :6], units, err_msg=f"{tr['map']} {tr['ai']} entry {k}")
            acts = np.asarray(e[4], np.int32).reshape(-1, 6)
            for p in (0, 1):
                r7 = np.asarray([[units[a[0]][2], units[a[0]][3], a[1], a[2], a[3], a[4], a[5]] for a in acts if units[a[0]][1] == p], np.int32).reshape(-1, 7)
                if len(r7):
                    _capi.check(L.mrts_debug_issue(h, 0, r7.ctypes.data_as(C.c_void_p), len(r7), None))
        L.mrts_destroy(h)


@pytest.mark.parametrize("mp,nsp,partial", [(MAP_8, 65536, False), (MAP_32, 2048, True)])
def test_full_size_properties(mp, nsp, partial):
    """BASELINE-sized slabs: size-independent properties (one plane per group per cell, mask rows
    only at source cells, NOOP always legal there, identical games stay identical under identical
    actions) plus oracle equality on a subsample of envs fed the same actions."""
    import torch

    env = _env(nsp, 0, partial_obs=partial, max_steps=64, map_paths=[mp], return_torch=True, seed=3)
    groups = 7 if partial else 6
    obs = env.reset()
    sub = 8
    ref = _oracle(sub, [], mp, partial, max_steps=64)
    np.testing.assert_array_equal(obs[:sub].cpu().numpy(), ref.reset())
    for t in range(100):
        m = env.get_action_mask()
        assert bool((obs.sum(3) == groups).all())
        src = m[:, :, 0] == 1
        assert bool((m.sum(2)[~src] == 0).all())
        a = env.sample_actions().clone()
        a[sub:2 * sub] = a[:sub]  # a second group of games fed the first group's actions
        o_ref, r_ref, d_ref = ref.step(a[:sub].cpu().numpy())
        obs, r, d, infos = env.step(a)
        np.testing.assert_array_equal(obs[:sub].cpu().numpy(), o_ref, err_msg=f"step {t}")
        np.testing.assert_array_equal(infos.raw[:sub].cpu().numpy(), r_ref)
        assert torch.equal(obs[:sub], obs[sub:2 * sub]) and torch.equal(d[:sub], d[sub:2 * sub])
    env.close()


@pytest.mark.parametrize("mp,n,partial,max_steps", [(MAP_8, 65536, False, 150), (MAP_32, 16384, True, 400)], ids=["cfg3_65536", "cfg4_16384"])
def test_baseline_sized_slabs_match_the_oracle_on_a_subsample(mp, n, partial, max_steps):
    """SURVEY 8(d) subsample protocol at BASELINE.json's sizes (configs[2]: 65,536 envs 8x8; configs[3]: 16,384 envs 32x32
    partial_obs): the whole slab steps on the device with device-sampled actions; 128 self-play games spread over the slab
    (256 envs) are mirrored on the oracle (all cores) with the same action rows for 1,000 steps, crossing max_steps resets."""
    import os

    import torch
    from oracle.oracle_py import OracleVecEnv

    sub_games = 128
    games = np.linspace(0, n // 2 - 1, sub_games).astype(np.int64)
    idx = np.stack([2 * games, 2 * games + 1], axis=1).reshape(-1)  # env rows of those games, pair by pair
    t_idx = torch.as_tensor(idx, device="cuda")
    env = _env(n, 0, partial_obs=partial, max_steps=max_steps, map_paths=[mp], return_torch=True, seed=11)
    ref = OracleVecEnv(len(idx), 0, max_steps, partial, templates_for([mp] * len(idx)), [], threads=os.cpu_count() or 1)
    obs = env.reset()
    np.testing.assert_array_equal(obs[t_idx].cpu().numpy(), ref.reset())
    n_done = 0
    for t in range(1000):
        m_ref, _ = ref.get_masks()  # every step: like the reference, the oracle reads the action rows of the cells its LAST mask flagged
        m = env.get_action_mask()
        if t % 25 == 0:
            np.testing.assert_array_equal(m[t_idx].cpu().numpy(), m_ref, err_msg=f"mask at step {t}")
        a = env.sample_actions()
        o_ref, r_ref, d_ref = ref.step(a[t_idx].cpu().numpy())
        obs, r, d, infos = env.step(a)
        np.testing.assert_array_equal(infos.raw[t_idx].cpu().numpy(), r_ref, err_msg=f"raw rewards at step {t}")
        np.testing.assert_array_equal(d[t_idx].cpu().numpy().astype(np.uint8), d_ref, err_msg=f"done at step {t}")
        if t % 5 == 0 or d_ref.any():
            np.testing.assert_array_equal(obs[t_idx].cpu().numpy(), o_ref, err_msg=f"observation at step {t}")
        n_done += int(d_ref.sum())
    assert n_done >= 2 * sub_games, "every mirrored game was meant to cross a reset"
    env.close()


@pytest.mark.parametrize("staged", [8, 12])
def test_games_outgrowing_the_staged_capacity_take_the_full_capacity_launch(staged):
    """The warp kernel stages only `staged_units` units per game in shared memory; games that may
    outgrow it are deferred to mrts_overflow_kernel (full blob capacity).  With a tiny staged
    capacity most WorkerRush games cross over (and come back after resets): every step must still
    equal the oracle bit for bit."""
    ais = [W, P, W, W]
    env = _env(4, 4, max_steps=400, ai2s=ais, map_paths=[MAP_16], staged_units=staged)
    ref = _oracle(4, ais, MAP_16, max_steps=400)
    rng = np.random.default_rng(21)
    np.testing.assert_array_equal(env.reset(), ref.reset())
    most = 0
    for t in range(900):
        m = env.get_action_mask()
        m_ref, _ = ref.get_masks()
        np.testing.assert_array_equal(m, m_ref, err_msg=f"mask {t}")
        a = sample_actions(m_ref, rng, 0.03)
        o, r, d, infos = env.step(a)
        o2, r2, d2 = ref.step(a)
        np.testing.assert_array_equal(o, o2, err_msg=f"obs {t}")
        np.testing.assert_array_equal(np.stack([i["raw_rewards"] for i in infos]), r2, err_msg=f"reward {t}")
        np.testing.assert_array_equal(d.astype(np.uint8), d2)
        if t % 50 == 0:
            most = max(most, max(len(env.export_state(g)[1]) for g in range(6)))
    assert most > max(staged, 8), "the case was meant to exceed the staged capacity"
    env.close()


def test_large_bot_slab_with_launch_ordering_matches_oracle():
    """Scripted-bot games are launched slowest class first (the pass appends every heavy game to one of four class lists
    for the next step, double-buffered; slot_game maps launch slots to games).  The permutation must not change any result: a subsample of a 4,096-env
    slab equals the oracle step by step, and every game is processed exactly once per step (the
    per-env step counters advance together: all dones fire on the same max_steps frame)."""
    import torch

    n, sub = 4096, 12
    ais = [W if i % 2 == 0 else P for i in range(n)]
    env = _env(0, n, max_steps=90, ai2s=ais, map_paths=[MAP_16], return_torch=True, seed=9)
    ref = _oracle(0, ais[:sub], MAP_16, max_steps=90)
    np.testing.assert_array_equal(env.reset()[:sub].cpu().numpy(), ref.reset())
    for t in range(200):
        m = env.get_action_mask()
        m_ref, _ = ref.get_masks()
        np.testing.assert_array_equal(m[:sub].cpu().numpy(), m_ref, err_msg=f"mask {t}")
        a = env.sample_actions()
        o, r, d, infos = env.step(a)
        o2, r2, d2 = ref.step(a[:sub].cpu().numpy())
        np.testing.assert_array_equal(o[:sub].cpu().numpy(), o2, err_msg=f"obs {t}")
        np.testing.assert_array_equal(infos.raw[:sub].cpu().numpy(), r2)
        if (t + 1) % 90 == 0:
            assert bool(d.all()), "every game must reach max_steps on the same frame"
        elif t < 60:
            assert not bool(d.any())
    env.close()


@pytest.mark.parametrize("mode", ["device", "host_transport"])
def test_cycle_maps_on_a_large_slab_follow_the_reference_order(mode):
    """Map rotation with more games than the rotation pass has threads (each thread then ranks a run of
    games) and every env ending on the same frame: the k-th ended env, ascending, must get the k-th next
    map of the shared list (vec_env.py:1038-1056), also through the packed host transport."""
    from tests.backends import make_env

    n = 2600
    ais = [W if i % 3 == 0 else P for i in range(n)]
    rotation = ["maps/8x8/basesWorkers8x8Obstacle.xml", "maps/8x8/FourBasesWorkers8x8.xml", MAP_8, "maps/8x8/melee8x8Mixed6.xml", "maps/8x8/basesWorkers8x8L.xml"]
    kw = dict(num_selfplay_envs=0, num_bot_envs=n, max_steps=23, ai2s=ais, map_paths=[MAP_8], cycle_maps=rotation)
    env = make_env("cuda", host_transport=(mode == "host_transport"), **kw)
    ref = make_env("oracle", **kw)
    rng = np.random.default_rng(5)
    np.testing.assert_array_equal(env.reset(), ref.reset())
    ended = 0
    for t in range(60):
        m_ref = ref.get_action_mask()
        np.testing.assert_array_equal(env.get_action_mask(), m_ref, err_msg=f"mask {t}")
        a = sample_actions(m_ref, rng, 0.02)
        o, r, d, _ = env.step(a)
        o2, r2, d2, _ = ref.step(a)
        np.testing.assert_array_equal(d, d2, err_msg=f"done {t}")
        np.testing.assert_array_equal(o, o2, err_msg=f"obs {t}")
        np.testing.assert_array_equal(r, r2)
        ended += int(d.sum())
    assert ended >= 2 * n
    env.close()


def test_step_into_writes_a_rollout_buffer_in_place():
    """step + masks written straight into slices of a preallocated rollout buffer
    (no intermediate tensors), next actions drawn on the device; every slice equals the oracle."""
    import torch

    ais = [W, P, W, LR]
    n, T = len(ais), 40
    env = _env(0, n, max_steps=25, ai2s=ais, map_paths=[MAP_8], return_torch=True, seed=4)
    ref = _oracle(0, ais, MAP_8, max_steps=25)
    dev = env.device
    obs = torch.zeros((T + 1, n, 8, 8, 29), dtype=torch.int32, device=dev)
    mask = torch.zeros((T + 1, n, 64, 78), dtype=torch.int32, device=dev)
    raw = torch.zeros((T, n, 6), dtype=torch.float64, device=dev)
    done = torch.zeros((T, n), dtype=torch.uint8, device=dev)
    act = torch.zeros((T + 1, n, 64, 7), dtype=torch.int32, device=dev)
    obs[0].copy_(env.reset())
    mask[0].copy_(env.get_action_mask())
    act[0].copy_(env.sample_actions())
    for t in range(T):
        env.step_into(act[t], obs[t + 1], mask[t + 1], raw[t], done[t], next_actions=act[t + 1])
    with pytest.raises(ValueError):
        env.step_into(act[0], obs[1][:, :4])  # wrong shape
    o, m, r, d, a = (x.cpu().numpy() for x in (obs, mask, raw, done, act))
    np.testing.assert_array_equal(o[0], ref.reset())
    for t in range(T):
        np.testing.assert_array_equal(m[t], ref.get_masks()[0], err_msg=f"mask {t}")
        o2, r2, d2 = ref.step(a[t])
        np.testing.assert_array_equal(o[t + 1], o2, err_msg=f"obs {t}")
        np.testing.assert_array_equal(r[t], r2)
        np.testing.assert_array_equal(d[t], d2)
    assert d.sum() >= n
    env.close()


def test_bot_vec_env_returns_rewards_and_dones_like_the_reference():
    """MicroRTSBotVecEnv (vec_env.py:1104-1236): np.ones observations, player 0's weighted reward, done,
    raw rewards in infos; the argument of step() is ignored."""
    from microrts_py_b200.envs import MicroRTSBotVecEnv
    from tests.backends import make_env

    ai1s, ai2s = [W, LR, P, LR], [LR, W, W, LR]
    rw = np.array([10.0, 1.0, 1.0, 0.2, 1.0, 4.0])
    env = MicroRTSBotVecEnv(ai1s=ai1s, ai2s=ai2s, max_steps=300, map_paths=[MAP_8], reward_weight=rw)
    ref = make_env("oracle", 0, 4, max_steps=300, ai2s=ai2s, ai1s=ai1s, map_paths=[MAP_8], reward_weight=rw)
    assert env.reset().shape == (4, 2) and env.observation_space.n == 2
    ref.reset()
    a = np.zeros((4, 64, 7), np.int32)
    ended = 0
    for t in range(700):
        o, r, d, infos = env.step(None)
        _, r2, d2, infos2 = ref.step(a)
        assert o.shape == (4, 2) and (o == 1).all()
        np.testing.assert_array_equal(r, r2, err_msg=f"reward {t}")
        np.testing.assert_array_equal(d, d2, err_msg=f"done {t}")
        np.testing.assert_array_equal(np.stack([i["raw_rewards"] for i in infos]), np.stack([i["raw_rewards"] for i in infos2]))
        ended += int(d.sum())
    assert ended >= 4
    env.close()
