#!/usr/bin/env python
"""Map an ncu SASS source page onto source lines (nvdisasm -g line info of the in-tree cubin).

    ncu -i X.ncu-rep --page source --csv > sass.csv
    python tools/ncu_lines.py sass.csv [kernel-substring] [top]

Prints the source lines with the most warp-stall samples and executed instructions."""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.environ.get("NCU_LINES_LIB") or os.path.join(ROOT, "microrts_py_b200", "csrc", "libmrts_b200.so")
SRC = os.environ.get("NCU_LINES_SRC") or os.path.join(ROOT, "microrts_py_b200", "csrc")
COL = os.environ.get("NCU_LINES_COL") or "# Samples"  # e.g. stall_long_sb, stall_no_inst


def line_table(kernel):
    d = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", LIB], cwd=d, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cubin)], capture_output=True, text=True).stdout
    table, cur, active, first = {}, ("?", 0), False, None
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            # one instantiation only: a substring that matches several template instances would merge their
            # offset tables -- give the mangled arguments too (e.g. mrts_pass_warp_kernelILi4ELi29ELi5E)
            active = kernel in ln and first in (None, ln)
            if active and first is None:
                first = ln
                print("section", ln.strip(), file=sys.stderr)
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*);", ln)
        if m:
            table[int(m.group(1), 16)] = (cur, m.group(2).strip())
    return table


def main():
    path = sys.argv[1]
    kernel = sys.argv[2] if len(sys.argv) > 2 else "mrts_rules_warp_kernel"
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    table = line_table(kernel)
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[h]
    ia, isamp, iinst = hdr.index("Address"), hdr.index(COL), hdr.index("Instructions Executed")
    body = [r for r in rows[h + 1 :] if len(r) > iinst and r[ia].startswith("0x")]
    base = int(body[0][ia], 16)
    per = defaultdict(lambda: [0, 0])
    tot = [0, 0]
    for r in body:
        off = int(r[ia], 16) - base
        key = table.get(off, (("?", 0), ""))[0]
        s, n = int(r[isamp] or 0), int(r[iinst] or 0)
        per[key][0] += s
        per[key][1] += n
        tot[0] += s
        tot[1] += n
    print(f"column {COL}: total {tot[0]}  warp instructions {tot[1]}")
    byfile = defaultdict(lambda: [0, 0])
    for (f, l), (s, n) in per.items():
        byfile[f][0] += s
        byfile[f][1] += n
    for f, (s, n) in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
        print(f"  {f:24s} samples {100*s/max(tot[0],1):5.1f}%  inst {100*n/max(tot[1],1):5.1f}%")
    src = {}
    for (f, l), (s, n) in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
        if f not in src:
            p = os.path.join(SRC, f)
            src[f] = open(p).read().splitlines() if os.path.exists(p) else []
        text = src[f][l - 1].strip() if 0 < l <= len(src[f]) else ""
        print(f"{100*s/max(tot[0],1):5.1f}% {100*n/max(tot[1],1):5.1f}%  {f}:{l}  {text[:110]}")


if __name__ == "__main__":
    main()
