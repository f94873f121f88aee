#!/bin/bash
mkdir -p gpurun_out
T=r02f
timeout 900 python -m pytest tests/test_trajectory_parity.py tests/test_import_state.py -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/${T}_summary.txt
tail -n 3 gpurun_out/${T}_pytest.log
B="python bench.py --steps 40 --warmup 5 --cpu-budget 0 --e2e-envs 0"
one() { tag=$1; shift
  env "$@" timeout 300 $B --workload $WL $EXTRA > gpurun_out/${T}_${WL}_$tag.json 2> gpurun_out/${T}_${WL}_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${T}_${WL}_$tag.json"))
    print("$WL $tag: %.2f M env-steps/s  %.3f ms  frac %.3f  launches %s %s" % (d["value"]/1e6, d["ms_per_step"], d["roofline"]["frac"], d["config"]["kernels_per_step"], d["config"]["rules_kernels"]))
except Exception as ex:
    print("$WL $tag: FAILED", ex)
PY
}
for WL in cfg2_rollout16 diag_wrush16 cfg1_selfplay16 cfg3_selfplay8 cfg4_partial32; do
  one auto MRTS_PIPELINE=1
done 2>&1 | tee -a gpurun_out/${T}_summary.txt
WL=cfg2_rollout16; EXTRA="--dtype uint8"; one u8 MRTS_PIPELINE=1 2>&1 | tee -a gpurun_out/${T}_summary.txt
EXTRA=""; N="python bench.py --steps 12 --warmup 2 --prewarm 200 --cpu-budget 0 --e2e-envs 0"
timeout 300 $N --workload diag_wrush16 > gpurun_out/${T}_plain1.log 2>&1 && \
timeout 1200 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --clock-control none --import-source on -k regex:mrts_rules_warp -s 205 -c 12 -f -o gpurun_out/${T}_wrush_src $N --workload diag_wrush16 > gpurun_out/${T}_ncu1.log 2>&1
ls -la gpurun_out/${T}*.ncu-rep
