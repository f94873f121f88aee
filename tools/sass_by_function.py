#!/usr/bin/env python
"""Instruction count of one kernel per source function (needs -lineinfo): which rule functions make up the
kernel image.  usage: sass_by_function.py <object or .so> <kernel name substring> [top]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

obj, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
d = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, stdout=subprocess.DEVNULL)
cub = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "--print-line-info", cub], capture_output=True, text=True).stdout.splitlines()
base = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "microrts_py_b200", "csrc")


def funcs(path):
    out = []
    for i, l in enumerate(open(path), 1):
        m = re.match(r"\s*(?:static\s+)?(?:MRTS_HDN?|__device__ __forceinline__|__global__)\s+[\w:<>\*&\s]*?\s\*?(\w+)\(", l)
        if m:
            out.append((i, m.group(1)))
    return out


fm = {f: funcs(os.path.join(base, f)) for f in os.listdir(base) if f.endswith((".h", ".cuh", ".cu"))}
inside, cur, cnt = False, None, collections.Counter()
for l in txt:
    if l.startswith(".text."):
        inside = pat in l
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l) and cur:
        cnt[cur] += 1
agg = collections.Counter()
for (f, ln), c in cnt.items():
    name = "?"
    for i, n in fm.get(f, []):
        if i <= ln:
            name = n
    agg[(f, name)] += c
tot = sum(agg.values())
print(f"{tot} instructions = {tot * 16 / 1024:.1f} KB")
for k, v in agg.most_common(top):
    print(f"{v:6d} {100 * v / tot:5.1f}%  {k[0]}:{k[1]}")
