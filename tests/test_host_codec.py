"""The PCIe transport codec of the host-buffer entry points (csrc/mrts_host.cpp), checked on CPU:
bit words -> one-hot arrays exactly as NumPy would unpack them, and action rows packed the way the
device decodes them (Sim::pack_action; UnitAction.fromVectorAction, J/rts/UnitAction.java:681-715).
The library is only loaded, no device call is made."""
import ctypes as C

import numpy as np
import pytest

from microrts_py_b200 import _capi


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.mark.parametrize("hw,planes,dtype", [(64, 29, np.int32), (256, 29, np.int32), (100, 31, np.int32), (81, 29, np.int32), (256, 31, np.uint8), (1, 29, np.int32)])
def test_observation_words_expand_to_one_hot(hw, planes, dtype):
    L = _capi.lib()
    rng = np.random.default_rng(hw * planes)
    n = 37
    words = rng.integers(0, 2 ** planes, size=(n, hw), dtype=np.uint32)
    want = ((words[:, :, None] >> np.arange(planes, dtype=np.uint32)) & 1).astype(dtype)
    for off in (0, 1):  # aligned and unaligned destination
        buf = np.zeros(n * hw * planes + 64, dtype)
        out = buf[off : off + n * hw * planes].reshape(n, hw, planes)
        _capi.check(L.mrts_host_expand_obs(_p(words), n, hw, planes, np.dtype(dtype).itemsize, _p(out)))
        np.testing.assert_array_equal(out, want)
        assert not buf[off + n * hw * planes :].any() and not buf[:off].any()


@pytest.mark.parametrize("dtype", [np.int32, np.uint8])
def test_mask_words_expand_to_78(dtype):
    L = _capi.lib()
    rng = np.random.default_rng(5)
    n, hw = 19, 144
    pm = np.zeros((n, hw, 3), np.uint32)
    cells = rng.random((n, hw)) < 0.05
    pm[cells] = rng.integers(0, 2 ** 32, size=(int(cells.sum()), 3), dtype=np.uint64).astype(np.uint32)
    pm[:, :, 2] &= (1 << 14) - 1  # 78 bits
    bits = np.concatenate([(pm[:, :, k : k + 1] >> np.arange(32, dtype=np.uint32)) & 1 for k in range(3)], axis=2)[:, :, :78]
    out = np.full((n, hw, 78), 7, dtype)
    _capi.check(L.mrts_host_expand_mask(_p(pm), n, hw, np.dtype(dtype).itemsize, _p(out)))
    np.testing.assert_array_equal(out, bits.astype(dtype))


def _pack_ref(a):
    t = int(a[0])
    if t < 1 or t > 5:
        return 0
    if t == 4:
        if not 0 <= a[5] < 7:
            return 0
        bad = not 0 <= a[4] <= 3
        pay = (0 if bad else int(a[4])) | (int(a[5]) << 2)
    elif t == 5:
        bad, pay = False, (63 if not 0 <= a[6] <= 48 else int(a[6]))
    else:
        bad = not 0 <= a[t] <= 3
        pay = 0 if bad else int(a[t])
    return t | (pay << 3) | (int(bad) << 9)


@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_action_rows_pack_only_at_source_cells(dtype):
    L = _capi.lib()
    rng = np.random.default_rng(11)
    n, hw = 23, 64
    a = rng.integers(-2, 9, size=(n, hw, 7)).astype(dtype)
    a[:, :, 6] = rng.integers(-3, 60, size=(n, hw))
    pm = np.zeros((n, hw, 3), np.uint32)
    src = rng.random((n, hw)) < 0.3
    pm[src, 0] = 1 | (rng.integers(0, 2 ** 20, size=int(src.sum())).astype(np.uint32) << 1)
    out = np.full((n, hw), 0xFFFF, np.uint16)
    _capi.check(L.mrts_host_pack_actions(_p(a), np.dtype(dtype).itemsize, _p(pm), n, hw, _p(out)))
    want = np.zeros((n, hw), np.uint16)
    for e, c in zip(*np.nonzero(src)):
        want[e, c] = _pack_ref(a[e, c])
    np.testing.assert_array_equal(out, want)


@pytest.mark.parametrize("hw,planes,dtype", [(256, 29, np.int32), (100, 31, np.int32), (64, 29, np.uint8), (3, 29, np.int32)])
def test_incremental_widening_tracks_a_changing_state(hw, planes, dtype):
    """mrts_*_host_slot: the shadow words say what the caller's array holds; after every update the array
    must equal the full expansion, nothing outside it may be touched, and an unchanged state writes nothing."""
    L = _capi.lib()
    rng = np.random.default_rng(planes * hw)
    n, elem = 23, np.dtype(dtype).itemsize
    words = rng.integers(0, 2 ** planes, size=(n, hw), dtype=np.uint32)
    pm = np.zeros((n, hw, 3), np.uint32)
    buf_o = np.zeros(n * hw * planes + 64, dtype)
    buf_m = np.zeros(n * hw * 78 + 64, dtype)
    out_o = buf_o[1 : 1 + n * hw * planes].reshape(n, hw, planes)
    out_m = buf_m[1 : 1 + n * hw * 78].reshape(n, hw, 78)
    _capi.check(L.mrts_host_expand_obs(_p(words), n, hw, planes, elem, _p(out_o)))
    _capi.check(L.mrts_host_expand_mask(_p(pm), n, hw, elem, _p(out_m)))
    sh_o, sh_m = words.copy(), pm.copy()
    for it in range(12):
        if it != 5:  # iteration 5: nothing changes
            k = rng.random((n, hw)) < (0.5 if it == 7 else 0.04)
            words[k] = rng.integers(0, 2 ** planes, size=int(k.sum()), dtype=np.uint32)
            pm[rng.random((n, hw)) < 0.03] = 0
            k = rng.random((n, hw)) < 0.03
            pm[k] = rng.integers(0, 2 ** 32, size=(int(k.sum()), 3), dtype=np.uint64).astype(np.uint32)
            pm[:, :, 2] &= (1 << 14) - 1
        _capi.check(L.mrts_host_expand_obs_delta(_p(words), _p(sh_o), n, hw, planes, elem, _p(out_o)))
        _capi.check(L.mrts_host_expand_mask_delta(_p(pm), _p(sh_m), n, hw, elem, _p(out_m)))
        np.testing.assert_array_equal(sh_o, words)
        np.testing.assert_array_equal(sh_m, pm)
        np.testing.assert_array_equal(out_o, ((words[:, :, None] >> np.arange(planes, dtype=np.uint32)) & 1).astype(dtype))
        bits = np.concatenate([(pm[:, :, j : j + 1] >> np.arange(32, dtype=np.uint32)) & 1 for j in range(3)], axis=2)[:, :, :78]
        np.testing.assert_array_equal(out_m, bits.astype(dtype))
        assert not buf_o[0] and not buf_o[1 + n * hw * planes :].any() and not buf_m[0] and not buf_m[1 + n * hw * 78 :].any()


@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_action_rows_pack_by_bitmap_touch_only_set_cells(dtype):
    """the device-delta slot path: source cells come as one bit per flat cell (csrc/mrts_delta.cu)"""
    L = _capi.lib()
    rng = np.random.default_rng(12)
    cells = 23 * 70 + 5  # not a multiple of 32
    a = rng.integers(-2, 9, size=(cells, 7)).astype(dtype)
    a[:, 6] = rng.integers(-3, 60, size=cells)
    src = rng.random(cells) < 0.2
    bits = np.zeros((cells + 31) // 32 * 32, np.uint8)
    bits[:cells] = src
    bitmap = np.packbits(bits.reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1)
    out = np.full(cells, 0xABCD, np.uint16)
    _capi.check(L.mrts_host_pack_actions_bitmap(_p(a), np.dtype(dtype).itemsize, _p(bitmap), cells, _p(out)))
    want = np.full(cells, 0xABCD, np.uint16)
    for c in np.nonzero(src)[0]:
        want[c] = _pack_ref(a[c])
    np.testing.assert_array_equal(out, want)


@pytest.mark.parametrize("dtype", [np.int32, np.int64])
def test_action_rows_pack_into_a_list_of_the_set_cells(dtype):
    """what the slot calls upload: {flat cell, packed row} of every set cell, each once, in any order; a short list is refused"""
    L = _capi.lib()
    rng = np.random.default_rng(13)
    cells = 50 * 64 + 11
    a = rng.integers(-2, 9, size=(cells, 7)).astype(dtype)
    a[:, 6] = rng.integers(-3, 60, size=cells)
    src = rng.random(cells) < 0.15
    bits = np.zeros((cells + 31) // 32 * 32, np.uint8)
    bits[:cells] = src
    bitmap = np.packbits(bits.reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1)
    n_src = int(src.sum())
    pairs = np.zeros((n_src + 5, 2), np.uint32)
    n = C.c_int64(0)
    _capi.check(L.mrts_host_pack_actions_sparse(_p(a), np.dtype(dtype).itemsize, _p(bitmap), cells, _p(pairs), len(pairs), C.byref(n)))
    assert n.value == n_src
    got = pairs[: n.value]
    order = np.argsort(got[:, 0])
    np.testing.assert_array_equal(got[order, 0], np.nonzero(src)[0])
    np.testing.assert_array_equal(got[order, 1], [_pack_ref(a[c]) for c in np.nonzero(src)[0]])
    assert not pairs[n.value:].any()
    short = np.zeros((10, 2), np.uint32)
    _capi.check(L.mrts_host_pack_actions_sparse(_p(a), np.dtype(dtype).itemsize, _p(bitmap), cells, _p(short), len(short), C.byref(n)))
    assert n.value == n_src > len(short)  # the caller sees that the list did not fit and uses the dense form


@pytest.mark.parametrize("planes,dtype", [(29, np.int32), (31, np.int32), (29, np.uint8), (31, np.uint8)])
def test_entries_patch_exactly_the_changed_range_of_their_cells(planes, dtype):
    """entries as mrts_delta_kernel emits them: new words + the first / last changed element per row; the arrays hold the
    expansion of the OLD words, after the patch they must hold the expansion of the new ones and nothing else may move"""
    L = _capi.lib()
    rng = np.random.default_rng(planes)
    cells, n = 4000, 700
    k78 = np.arange(78)

    def expand(ow, mw):
        o = ((ow[:, None] >> np.arange(planes, dtype=np.uint32)) & 1).astype(dtype)
        m = ((mw[:, k78 // 32] >> (k78 % 32).astype(np.uint32)) & 1).astype(dtype)
        return np.ascontiguousarray(o), np.ascontiguousarray(m)  # (fancy indexing may hand back another memory order)

    old_o = rng.integers(0, 2 ** planes, size=cells, dtype=np.uint64).astype(np.uint32)
    old_m = rng.integers(0, 2 ** 32, size=(cells, 3), dtype=np.uint64).astype(np.uint32)
    old_m[:, 2] &= (1 << 14) - 1
    obs, mask = expand(old_o, old_m)
    picked = rng.choice(cells, size=n, replace=False)
    new_o, new_m = old_o.copy(), old_m.copy()
    flips = rng.integers(0, 2 ** planes, size=n, dtype=np.uint64).astype(np.uint32)
    flips[::5] = 0            # observation row unchanged
    flips[1::7] = 1 << 3      # a single plane
    new_o[picked] ^= flips
    mflip = rng.integers(0, 2 ** 32, size=(n, 3), dtype=np.uint64).astype(np.uint32)
    mflip[:, 2] &= (1 << 14) - 1
    mflip[::3] = 0            # mask row unchanged
    mflip[1::4, 0] = 0        # change starts in the second word
    mflip[2::9, 1:] = 0       # change confined to the first word
    mflip[5::5] = 0
    mflip[5::5, 1] = 1 << 9   # a single entry
    new_m[picked] ^= mflip
    both_same = (flips == 0) & (mflip == 0).all(axis=1)
    ent = np.zeros((n, 6), np.uint32)
    ent[:, 0] = picked
    ent[:, 1] = new_o[picked]
    ent[:, 2:5] = new_m[picked]
    for i in range(n):
        xo = int(flips[i])
        ko0, ko1 = (31, 0) if xo == 0 else ((xo & -xo).bit_length() - 1, xo.bit_length() - 1)
        xm = int(mflip[i, 0]) | (int(mflip[i, 1]) << 32) | (int(mflip[i, 2]) << 64)
        kb0, kb1 = (127, 0) if xm == 0 else ((xm & -xm).bit_length() - 1, xm.bit_length() - 1)
        ent[i, 5] = ko0 | (ko1 << 5) | (kb0 << 10) | (kb1 << 17)
    ent = np.ascontiguousarray(ent[~both_same])
    want_obs, want_mask = expand(new_o, new_m)
    _capi.check(L.mrts_host_patch_cells(_p(ent), len(ent), planes, np.dtype(dtype).itemsize, _p(obs), _p(mask)))
    np.testing.assert_array_equal(obs, want_obs)
    np.testing.assert_array_equal(mask, want_mask)
    _capi.check(L.mrts_host_patch_cells(_p(ent), len(ent), planes, np.dtype(dtype).itemsize, _p(obs), None))  # either array may be null
