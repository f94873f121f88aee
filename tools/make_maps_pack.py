#!/usr/bin/env python
"""Generate microrts_py_b200/data/maps_pack.json.gz from the reference's maps/ tree.

Run in the build container (the reference is mounted at /root/reference):
    python tools/make_maps_pack.py [/root/reference/gym_microrts/microrts]
Every ``maps/**.xml`` is parsed by microrts_py_b200.maps.parse_map_xml and stored as a flat
template keyed by its path relative to the microrts root (the string users pass in
``map_paths``).  The GPU box has no /root/reference, so the pack is committed.
"""
import gzip
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from microrts_py_b200.maps import parse_map_xml  # noqa: E402


def main():
    root = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/gym_microrts/microrts"
    out = {}
    for d, _, files in os.walk(os.path.join(root, "maps")):
        for f in sorted(files):
            if not f.endswith(".xml"):
                continue
            p = os.path.join(d, f)
            rel = os.path.relpath(p, root).replace(os.sep, "/")
            try:
                out[rel] = parse_map_xml(p, rel).to_json()
            except Exception as e:  # noqa: BLE001
                print("skip", rel, e)
    dst = os.path.join(os.path.dirname(__file__), "..", "microrts_py_b200", "data", "maps_pack.json.gz")
    with gzip.GzipFile(dst, "wb", mtime=0) as g:
        g.write(json.dumps(out, sort_keys=True, separators=(",", ":")).encode())
    print(f"{len(out)} maps -> {dst} ({os.path.getsize(dst)} bytes)")


if __name__ == "__main__":
    main()
