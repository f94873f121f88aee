#!/usr/bin/env python
"""DRAM traffic per step from an `ncu --set full` capture -> profiles/traffic_<workload>_<dtype>.json (read by bench.py).

  python tools/ncu_traffic.py gpurun_out/r02g_full_cfg2_rollout16.ncu-rep --workload cfg2_rollout16 --dtype int32 --steps 2

The capture holds the kernels of `--steps` whole steps (tools/gpu_measure.sh `full`: -c = launches per step x steps).
dram__bytes_read.sum / dram__bytes_write.sum are summed over the captured launches and divided by the step count; the
file records the kernel sources the capture was taken on (bench.source_hash) so that bench.py refuses a stale one.
ncu replays each kernel with cold caches: the figure is an upper bound on a warm step's traffic."""
import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

UNITS = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out[out.index('"ID"'):])))
    head, units, body = rows[0], rows[1], rows[2:]
    return head, units, body


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--workload", required=True)
    ap.add_argument("--dtype", default="int32")
    ap.add_argument("--steps", type=int, required=True, help="whole steps the capture covers")
    ap.add_argument("--envs", type=int, default=0, help="envs of the captured run (default: the workload's)")
    ap.add_argument("--commit", default="")
    a = ap.parse_args()
    head, units, body = raw_rows(a.rep)
    col = {h: i for i, h in enumerate(head)}
    name_i = col["Kernel Name"]

    def val(row, metric):
        i = col[metric]
        return float(row[i].replace(",", "")) * UNITS.get(units[i], 1)

    rd = sum(val(r, "dram__bytes_read.sum") for r in body)
    wr = sum(val(r, "dram__bytes_write.sum") for r in body)
    dur = sum(val(r, "gpu__time_duration.sum") for r in body) if "gpu__time_duration.sum" in col else None
    kernels = {}
    for r in body:
        k = r[name_i].split("(")[0]
        kernels[k] = kernels.get(k, 0) + 1
    wl = bench.workload_setup(a.workload, a.envs or None)
    n, hw, planes = wl["n"], bench.CELLS[wl["map"]], wl["planes"]
    elem = 4 if a.dtype == "int32" else 1
    B = bench.algorithmic_bytes(hw, planes, elem)
    commit = a.commit or subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    t = {
        "workload": a.workload, "dtype": a.dtype, "envs": n, "steps_captured": a.steps, "launches_captured": len(body), "kernels": kernels,
        "dram_read_bytes": rd / a.steps, "dram_write_bytes": wr / a.steps, "dram_bytes_per_step": (rd + wr) / a.steps,
        "algorithmic_bytes_per_step": B * n, "dram_over_algorithmic": (rd + wr) / a.steps / (B * n),
        "kernel_time_ns_per_step_under_ncu": dur / a.steps if dur else None,
        "source_hash": bench.source_hash(), "commit": commit, "capture": os.path.basename(a.rep),
        "note": "ncu --set full --clock-control none; every kernel replayed with cold caches",
    }
    p = os.path.join(ROOT, "profiles", f"traffic_{a.workload}_{a.dtype}.json")
    json.dump(t, open(p, "w"), indent=1)
    print(json.dumps(t, indent=1))
    print("wrote", p)


if __name__ == "__main__":
    main()
