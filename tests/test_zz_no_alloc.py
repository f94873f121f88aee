"""mrts_expand_static: the allocation-free, env-less expansion entry point (include/mrts_b200.h) against mrts_expand.

The file is named so that pytest collects it LAST.  The entry point has no counterpart in the reference:
its scratch memory is a file-scope __device__ array whose device address the library
takes with cudaGetSymbolAddress, so the checks are (1) bit-exact equality with the existing entry point on the same
packed inputs, on two calls in a row that share the scratch array, (2) every output byte is written (poisoned buffers),
(3) the size limit is refused with a clear error, (4) the calls leave the free device memory unchanged."""
import ctypes as C

import numpy as np
import pytest

from microrts_py_b200 import _capi, microrts_ai
from tests.common import MAP_8, MAP_16, sample_actions

MAX_VIEWS = 65536  # MRTS_EXPAND_STATIC_MAX_VIEWS


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def test_rejects_more_views_than_the_static_scratch_holds():
    """refused before any CUDA call: runs without a GPU"""
    L = _capi.lib()
    rc = L.mrts_expand_static(C.c_void_p(16), C.c_void_p(16), C.c_void_p(16), C.c_void_p(16), MAX_VIEWS + 1, 16, 16, 64, 0, 4, None)
    assert rc == -1
    msg = L.mrts_last_error().decode()
    assert "MRTS_EXPAND_STATIC_MAX_VIEWS" in msg and str(MAX_VIEWS + 1) in msg
    assert L.mrts_expand_static(C.c_void_p(16), C.c_void_p(16), C.c_void_p(16), C.c_void_p(16), 0, 16, 16, 64, 0, 4, None) == -1
    # 15 x 14 cells of int32: view blocks are not multiples of 16 bytes (mrts_expand refuses the same shape)
    assert L.mrts_expand_static(C.c_void_p(16), C.c_void_p(16), C.c_void_p(16), C.c_void_p(16), 4, 14, 15, 64, 0, 4, None) == -1
    assert "multiple of 16" in L.mrts_last_error().decode()


CASES = {
    "i32_p29_16x16": (np.int32, False, MAP_16, 24),
    "u8_p31_8x8": (np.uint8, True, MAP_8, 40),
    "i32_p31_16x16": (np.int32, True, MAP_16, 9),
}


@pytest.mark.gpu
@pytest.mark.parametrize("case", list(CASES))
def test_static_expand_matches_mrts_expand_on_two_calls_in_a_row(case):
    import torch

    from microrts_py_b200.envs import MicroRTSGridModeVecEnv

    dtype, partial, mp, n_bot = CASES[case]
    W, LR, P = microrts_ai.workerRushAI, microrts_ai.lightRushAI, microrts_ai.passiveAI
    env = MicroRTSGridModeVecEnv(4, n_bot, partial_obs=partial, max_steps=300, ai2s=[[W, LR, P][k % 3] for k in range(n_bot)], map_paths=[mp] * (4 + n_bot),
                                 reward_weight=np.array([10.0, 1.0, 1.0, 0.2, 1.0, 4.0]), dtype=dtype, return_packed=True)
    L, n, h, w = env._lib, env.num_envs, env.height, env.width
    cap = L.mrts_unit_capacity(env._h)
    planes = 31 if partial else 29
    elem = np.dtype(dtype).itemsize
    t_dtype = torch.int32 if elem == 4 else torch.uint8
    rng = np.random.default_rng(5)

    # two different packed snapshots: after reset + 30 steps, and 40 steps later
    snaps = []
    words = env.reset()
    for t in range(70):
        entries = env.get_action_mask()
        if t in (30, 69):
            snaps.append((words.clone(), entries.clone()))
        mask = env.expand(None, entries)[1]
        words, _, _, _ = env.step(sample_actions(mask.cpu().numpy().astype(np.int32), rng, 0.03))
    assert not torch.equal(snaps[0][0], snaps[1][0])
    want = [env.expand(wd, en) for wd, en in snaps]  # the existing entry point
    torch.cuda.synchronize()

    def poisoned():
        return (torch.full((n, h, w, planes), 0x5A, dtype=t_dtype, device=words.device), torch.full((n, h * w, 78), 0x5A, dtype=t_dtype, device=words.device))

    def static_expand(wd, en, obs, mask):
        with torch.cuda.device(words.device):
            rc = L.mrts_expand_static(_p(wd), _p(en), _p(obs), _p(mask), n, h, w, cap, int(partial), elem, env._stream())
        assert rc == 0, L.mrts_last_error().decode()

    got = [poisoned(), poisoned()]
    static_expand(*snaps[0], *got[0])  # first call on this device in a fresh process: does the one-time set-up
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(words.device)[0]
    static_expand(*snaps[1], *got[1])  # two calls in a row, same stream, same scratch
    static_expand(*snaps[0], *got[0])
    static_expand(*snaps[1], *got[1])
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(words.device)[0] >= free_before, "the entry point must not allocate device memory"
    for k in range(2):
        assert torch.equal(got[k][0], want[k][0]), f"observation of call {k}"
        assert torch.equal(got[k][1], want[k][1]), f"mask of call {k}"

    # one tensor at a time (the other pointer NULL), again back to back
    o_only, m_only = poisoned()[0], poisoned()[1]
    static_expand(snaps[0][0], None, o_only, None)
    static_expand(None, snaps[1][1], None, m_only)
    torch.cuda.synchronize()
    assert torch.equal(o_only, want[0][0]) and torch.equal(m_only, want[1][1])
    env.close()
