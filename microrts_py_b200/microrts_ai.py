"""Scripted opponents usable as ``ai2s`` entries.

Mirrors gym_microrts/microrts_ai.py: each name is a callable taking the unit type table and
returning the opponent.  In the reference the result is a Java ``ai.core.AI`` object
(microrts_ai.py:13-22); here it is a small descriptor naming the device-side implementation
(``MRTS_AI_*`` in include/mrts_b200.h).  Only the deterministic bots the hot path covers exist:

  passiveAI     ai.PassiveAI             (J/ai/PassiveAI.java:41-45)
  workerRushAI  ai.abstraction.WorkerRush (J/ai/abstraction/WorkerRush.java:63-204)
  lightRushAI   ai.abstraction.LightRush  (J/ai/abstraction/LightRush.java:77-254)
  POWorkerRush / POLightRush  the same two with the exploring melee behaviour of
                J/ai/abstraction/partialobservability/PO{Worker,Light}Rush.java:42-78
  POHeavyRush / PORangedRush  POLightRush training heavy / ranged units (J/ai/abstraction/HeavyRush.java:136-137,
                RangedRush.java:124-125 are the only lines that differ from LightRush.java)

Every other factory of the reference (random, search-based or closed-jar bots) raises
``NotImplementedError`` with the reason.
"""
from __future__ import annotations


class DeviceAI:
    def __init__(self, name: str, ai_id: int):
        self.name = name
        self.ai_id = ai_id

    def __repr__(self):
        return f"DeviceAI({self.name})"


def passiveAI(utt=None):
    return DeviceAI("PassiveAI", 0)


def workerRushAI(utt=None):
    return DeviceAI("WorkerRush", 1)


def lightRushAI(utt=None):
    return DeviceAI("LightRush", 2)


def POWorkerRush(utt=None):
    return DeviceAI("POWorkerRush", 3)


def POLightRush(utt=None):
    return DeviceAI("POLightRush", 4)


def POHeavyRush(utt=None):
    return DeviceAI("POHeavyRush", 5)


def PORangedRush(utt=None):
    return DeviceAI("PORangedRush", 6)


def _unsupported(name, why):
    def f(utt=None):
        raise NotImplementedError(f"{name}: {why}")

    f.__name__ = name
    return f


_RANDOM = "draws from an unseeded java.util.Random; not reproducible, outside the device hot path"
_SEARCH = "wall-clock-budgeted search / closed competition jar; outside the device hot path"
randomBiasedAI = _unsupported("randomBiasedAI", _RANDOM)
randomAI = _unsupported("randomAI", _RANDOM)
for _n in ("naiveMCTSAI", "mixedBot", "rojo", "izanagi", "tiamat", "droplet", "guidedRojoA3N", "coacAI", "mayari"):
    globals()[_n] = _unsupported(_n, _SEARCH)
del _n

ALL_AIS = [passiveAI, workerRushAI]
