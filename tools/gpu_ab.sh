#!/bin/bash
# A/B of the pipelines on one B200: direct vs split with 1/2/4/8 chunks, device-resident numbers only.
# usage: tools/gpu_ab.sh <tag> [workloads...]
tag=${1:-ab}; shift
wls=${@:-cfg2_rollout16 diag_wrush16 diag_passive16 cfg1_selfplay16 cfg3_selfplay8}
mkdir -p gpurun_out
for wl in $wls; do
  for cfgs in "1 0" "2 1" "2 2" "2 4" "2 8"; do
    set -- $cfgs
    MRTS_PIPELINE=$1 MRTS_CHUNKS=$2 timeout 300 python bench.py --workload $wl --steps 60 --warmup 5 --cpu-budget 0 --e2e-envs 0 \
      > gpurun_out/${tag}_${wl}_p$1_k$2.json 2> gpurun_out/${tag}_${wl}_p$1_k$2.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${tag}_${wl}_p$1_k$2.json"))
    print("$wl pipeline=$1 chunks=$2: %.2f M env-steps/s  %.3f ms  frac %.3f  launches %s" % (d["value"]/1e6, d["ms_per_step"], d["roofline"]["frac"], d["config"]["kernels_per_step"]))
except Exception as ex:
    print("$wl pipeline=$1 chunks=$2: FAILED", ex)
PY
  done
done
