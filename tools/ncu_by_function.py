#!/usr/bin/env python
"""Stall samples / executed instructions of one ncu capture, summed per source function and per source line.
    ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv ; python tools/ncu_by_function.py src.csv [top]
Needs a capture taken with --import-source on (the CUDA-C view carries file + line)."""
import collections
import csv
import os
import re
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
base = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "microrts_py_b200", "csrc")


def funcs(p):
    out = []
    for i, l in enumerate(open(p), 1):
        m = re.match(r"\s*(?:template\s*<[^>]*>\s*)?(?:static\s+)?(?:MRTS_HDN?|__device__ __forceinline__|__global__)\s+[\w:<>\*&\s,]*?\s\*?(\w+)\(", l)
        if m:
            out.append((i, m.group(1)))
    return out


fm = {f: funcs(os.path.join(base, f)) for f in os.listdir(base) if f.endswith((".h", ".cuh", ".cu"))}
rows = list(csv.reader(open(path)))
cur_file, hdr = None, None
STALLS = ["stall_long_sb", "stall_short_sb", "stall_wait", "stall_no_inst", "stall_lg", "stall_mio", "stall_barrier", "stall_branch_resolving", "stall_math", "stall_dispatch", "stall_not_selected", "stall_selected", "stall_drain", "stall_membar", "stall_sleep", "stall_misc"]
per_fn = collections.defaultdict(lambda: collections.Counter())
per_line = collections.defaultdict(lambda: collections.Counter())
tot = collections.Counter()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = os.path.basename(r[1])
        continue
    if r[0] == "Line No":
        hdr = r
        idx = {n: i for i, n in enumerate(hdr)}
        continue
    if hdr is None or cur_file is None or len(r) < len(hdr) or not r[0].isdigit():
        continue
    ln = int(r[0])
    name = "?"
    for i, n in fm.get(cur_file, []):
        if i <= ln:
            name = n
    def col(n):  # a source line with unescaped quotes splits into extra fields: count columns from the end
        v = r[idx[n] - len(hdr)]
        try:
            return int(float(v or 0))
        except ValueError:
            return 0

    vals = {"samples": col("# Samples"), "inst": col("Instructions Executed")}
    for s in STALLS:
        if s in idx:
            vals[s] = col(s)
    key = (cur_file, name)
    for k, v in vals.items():
        per_fn[key][k] += v
        per_line[(cur_file, ln, r[1].strip()[:90])][k] += v
        tot[k] += v
print("total samples", tot["samples"], "instructions", tot["inst"])
print("stall mix: " + ", ".join(f"{s[6:]} {100 * tot[s] / max(tot['samples'], 1):.1f}%" for s in sorted(STALLS, key=lambda s: -tot[s])[:8]))
print("\n-- per function (samples %, instructions %, top stalls)")
for key, c in sorted(per_fn.items(), key=lambda kv: -kv[1]["samples"])[:top]:
    st = ", ".join(f"{s[6:]} {100 * c[s] / max(c['samples'], 1):.0f}%" for s in sorted(STALLS, key=lambda s: -c[s])[:3])
    print(f"{100 * c['samples'] / max(tot['samples'], 1):5.1f}%  {100 * c['inst'] / max(tot['inst'], 1):5.1f}%  {key[0]}:{key[1]:28s} {st}")
print("\n-- per line")
for key, c in sorted(per_line.items(), key=lambda kv: -kv[1]["samples"])[:top]:
    st = ", ".join(f"{s[6:]} {100 * c[s] / max(c['samples'], 1):.0f}%" for s in sorted(STALLS, key=lambda s: -c[s])[:2])
    print(f"{100 * c['samples'] / max(tot['samples'], 1):5.1f}%  {100 * c['inst'] / max(tot['inst'], 1):5.1f}%  {key[0]}:{key[1]:5d}  {st:32s} {key[2]}")
