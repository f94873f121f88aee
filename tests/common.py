"""Shared helpers of the test-suite: the seeded masked-uniform action sampler (the recipe of the
reference's hello_world.py:27-67: an independent categorical per action component restricted to
mask == 1; an all-zero group samples uniformly over the whole group) and config builders."""
import numpy as np

from microrts_py_b200.maps import load_map

GROUPS = [(0, 6), (6, 10), (10, 14), (14, 18), (18, 22), (22, 29), (29, 78)]

MAP_8 = "maps/8x8/basesWorkers8x8.xml"
MAP_16 = "maps/16x16/basesWorkers16x16A.xml"
MAP_32 = "maps/BWDistantResources32x32.xml"
MAP_64 = "maps/BroodWar/(4)BloodBath.scmB.xml"
MAP_4 = "maps/4x4/baseTwoWorkers4x4.xml"


def sample_actions(mask, rng, eps=0.0):
    """mask (N, HW, 78) of 0/1 -> actions (N, HW, 7) int32.  With probability `eps` a cell's
    components are drawn ignoring the mask (exercises the illegal-action rewrite)."""
    n, hw, _ = mask.shape
    out = np.zeros((n, hw, 7), np.int32)
    wild = rng.random((n, hw)) < eps if eps > 0 else None
    for k, (a, b) in enumerate(GROUPS):
        m = mask[:, :, a:b].astype(np.float64)
        if wild is not None:
            m = np.where(wild[:, :, None], 0.0, m)
        score = rng.random((n, hw, b - a)) + 10.0 * m
        out[:, :, k] = score.argmax(axis=2)
    return out


def templates_for(paths):
    cache = {}
    out = []
    for p in paths:
        if p not in cache:
            cache[p] = load_map(p)
        out.append(cache[p])
    return out
