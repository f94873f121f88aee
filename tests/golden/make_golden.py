#!/usr/bin/env python
"""Convert the reference's 280 golden game traces into one compact fixture.

Run in the build container only (the reference is mounted at /root/reference):
    python tests/golden/make_golden.py
Source: gym_microrts/microrts/data/traces/<map path minus .xml>/<AI>/trace_0.zip, written by
gym_microrts/microrts/src/tests/GenerateTestTraces.java:102-134 (each zip holds one member whose
bytes are the XML of an rts.Trace).  Output: tests/golden/traces.json.gz with, per trace,
  {"map": "maps/....xml", "ai": "LightRush"|"PortfolioAI", "entries": [[time, res0, res1,
    [type,player,x,y,resources,hitpoints]*U (flattened), [unit_index,type,param,x,y,utype]*A (flattened)], ...]}
Actions reference their unit by its index in the same entry's unit list (trace unit IDs are
arbitrary); utype is the produced type id or -1; a missing XML `parameter` is stored as -1
(UnitAction's field default, rts/UnitAction.java:114).
"""
import gzip
import json
import os
import sys
import xml.etree.ElementTree as ET
import zipfile

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from microrts_py_b200.maps import UNIT_TYPE_NAMES  # noqa: E402

TID = {n: i for i, n in enumerate(UNIT_TYPE_NAMES)}


def convert(path):
    z = zipfile.ZipFile(path)
    root = ET.fromstring(z.read(z.namelist()[0]))
    entries = []
    for te in root.find("entries"):
        t = int(te.get("time"))
        pgs = te.find("rts.PhysicalGameState")
        res = [0, 0]
        for p in pgs.find("players"):
            res[int(p.get("ID"))] = int(p.get("resources"))
        units, idx = [], {}
        for k, u in enumerate(pgs.find("units")):
            idx[u.get("ID")] = k
            units += [TID[u.get("type")], int(u.get("player")), int(u.get("x")), int(u.get("y")), int(u.get("resources")), int(u.get("hitpoints"))]
        acts = []
        for a in te.find("actions"):
            ua = a.find("UnitAction")
            ut = ua.get("unitType")
            acts += [
                idx[a.get("unitID")],
                int(ua.get("type")),
                int(ua.get("parameter")) if ua.get("parameter") is not None else -1,
                int(ua.get("x")) if ua.get("x") is not None else 0,
                int(ua.get("y")) if ua.get("y") is not None else 0,
                TID[ut] if ut is not None else -1,
            ]
        entries.append([t, res[0], res[1], units, acts])
    return entries


def main():
    root = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/gym_microrts/microrts"
    troot = os.path.join(root, "data", "traces")
    out = []
    for d, _, files in sorted(os.walk(troot)):
        for f in files:
            if f != "trace_0.zip":
                continue
            rel = os.path.relpath(d, troot).replace(os.sep, "/")
            mapname, ai = rel.rsplit("/", 1)
            out.append({"map": "maps/" + mapname + ".xml", "ai": ai.split("_")[0], "entries": convert(os.path.join(d, f))})
    out.sort(key=lambda e: (e["map"], e["ai"]))
    dst = os.path.join(os.path.dirname(__file__), "traces.json.gz")
    with gzip.GzipFile(dst, "wb", mtime=0) as g:
        g.write(json.dumps(out, separators=(",", ":")).encode())
    print(f"{len(out)} traces -> {dst} ({os.path.getsize(dst)} bytes)")


if __name__ == "__main__":
    main()
