"""Map templates: the host-side loader for microRTS map XML files.

Replaces ``PhysicalGameState.load/fromXML`` (gym_microrts/microrts/src/rts/PhysicalGameState.java:
54-65,685-711,750-762), ``Unit.fromXML`` (rts/units/Unit.java:597-620) and ``Player.fromXML``
(rts/Player.java:95-99).  The reference re-parses the XML on every episode reset
(tests/JNIGridnetClient.java:232); here a map is parsed once into a flat template that is
uploaded to the device and every reset is a device-side copy.

``map_paths`` entries are resolved like the reference does (relative to the microrts root,
e.g. ``maps/16x16/basesWorkers16x16A.xml``): an existing file (absolute, relative to the cwd or
to ``microrts_path``) is parsed; otherwise the name is looked up in the bundled map pack
(``data/maps_pack.json.gz``, generated from the reference's ``maps/`` tree by
``tools/make_maps_pack.py``), so the GPU box needs no checkout of the reference.
"""
from __future__ import annotations

import gzip
import json
import os
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

UNIT_TYPE_NAMES = ["Resource", "Base", "Barracks", "Worker", "Light", "Heavy", "Ranged"]
_TYPE_ID = {n: i for i, n in enumerate(UNIT_TYPE_NAMES)}

_PACK_PATH = os.path.join(os.path.dirname(__file__), "data", "maps_pack.json.gz")
_pack_cache: Optional[Dict[str, dict]] = None


@dataclass
class MapTemplate:
    """Flat map template. ``units`` rows are (id, type, player, x, y, resources, hitpoints) in XML
    order, which is the initial unit order (``pgs.units`` is a LinkedList in creation order)."""

    name: str
    width: int
    height: int
    terrain: np.ndarray  # (H*W,) uint8, row-major, 0 free / 1 wall
    resources: List[int] = field(default_factory=lambda: [0, 0])
    units: np.ndarray = field(default_factory=lambda: np.zeros((0, 7), np.int32))

    def to_json(self) -> dict:
        return {
            "w": self.width,
            "h": self.height,
            "t": "".join(str(int(v)) for v in self.terrain),
            "r": list(map(int, self.resources)),
            "u": self.units.astype(int).reshape(-1).tolist(),
        }

    @staticmethod
    def from_json(name: str, d: dict) -> "MapTemplate":
        return MapTemplate(
            name=name,
            width=int(d["w"]),
            height=int(d["h"]),
            terrain=np.frombuffer(d["t"].encode(), np.uint8) - ord("0"),
            resources=list(d["r"]),
            units=np.asarray(d["u"], np.int32).reshape(-1, 7),
        )


def _uncompress_terrain(t: str) -> List[int]:
    """Run-length ``A``/``B`` terrain form, PhysicalGameState.java:562-592."""
    out: List[int] = []
    counter = ""
    for ch in t:
        if ch in "AB":
            if counter:
                out.extend([out[-1]] * (int(counter) - 1))
                counter = ""
            out.append(0 if ch == "A" else 1)
        else:
            counter += ch
    if counter:
        out.extend([out[-1]] * (int(counter) - 1))
    return out


def parse_map_xml(text_or_path: str, name: str = "") -> MapTemplate:
    if os.path.exists(text_or_path):
        root = ET.parse(text_or_path).getroot()
        name = name or text_or_path
    else:
        root = ET.fromstring(text_or_path)
    w, h = int(root.get("width")), int(root.get("height"))
    tstr = root.find("terrain").text or ""
    if "A" in tstr or "B" in tstr:
        terr = _uncompress_terrain(tstr)
    else:
        terr = [int(c) for c in tstr[: w * h]]  # getTerrainFromUnknownString reads exactly w*h chars
    if len(terr) != w * h:
        raise ValueError(f"map {name}: terrain has {len(terr)} cells, expected {w * h}")
    res = [0, 0]
    for p in root.find("players"):
        res[int(p.get("ID"))] = int(p.get("resources"))
    rows = []
    seen = set()
    for u in root.find("units"):
        uid = int(u.get("ID"))
        if uid in seen:
            raise ValueError(f"Repeated unit ID {uid} in map!")  # PhysicalGameState.java:702-704
        seen.add(uid)
        tname = u.get("type")
        if tname not in _TYPE_ID:
            raise ValueError(f"unknown unit type {tname!r} in map {name}")
        rows.append(
            [uid, _TYPE_ID[tname], int(u.get("player")), int(u.get("x")), int(u.get("y")), int(u.get("resources")), int(u.get("hitpoints"))]
        )
    return MapTemplate(name, w, h, np.asarray(terr, np.uint8), res, np.asarray(rows, np.int32).reshape(-1, 7))


def _pack() -> Dict[str, dict]:
    global _pack_cache
    if _pack_cache is None:
        if os.path.exists(_PACK_PATH):
            with gzip.open(_PACK_PATH, "rt") as f:
                _pack_cache = json.load(f)
        else:
            _pack_cache = {}
    return _pack_cache


def bundled_map_names() -> List[str]:
    return sorted(_pack().keys())


def load_map(path: str, microrts_path: Optional[str] = None) -> MapTemplate:
    """Resolve and load one ``map_paths`` entry."""
    cands = [path]
    if microrts_path:
        cands.append(os.path.join(microrts_path, path))
    for c in cands:
        if os.path.isfile(c):
            return parse_map_xml(c, name=path)
    key = path.replace("\\", "/")
    if key.startswith("./"):
        key = key[2:]
    pk = _pack()
    if key in pk:
        return MapTemplate.from_json(key, pk[key])
    # also accept a path that merely ends with a bundled name
    for k in pk:
        if key.endswith(k):
            return MapTemplate.from_json(k, pk[k])
    raise FileNotFoundError(f"map {path!r} is neither a file nor in the bundled map pack")
