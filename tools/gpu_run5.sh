#!/bin/bash
mkdir -p gpurun_out
T=r02e
timeout 1200 python -m pytest tests/test_trajectory_parity.py tests/test_gpu_api.py tests/test_known_answers.py -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest default rc=$?" | tee gpurun_out/${T}_summary.txt
timeout 1500 python -m pytest tests/test_parity_matrix.py tests/test_reference_consumers.py -m gpu -x -q > gpurun_out/${T}_pytest_matrix.log 2>&1; echo "pytest matrix rc=$?" | tee -a gpurun_out/${T}_summary.txt
tail -n 3 gpurun_out/${T}_pytest.log gpurun_out/${T}_pytest_matrix.log
B="python bench.py --steps 40 --warmup 5 --cpu-budget 0 --e2e-envs 0"
one() { tag=$1; shift
  env "$@" timeout 300 $B --workload $WL > gpurun_out/${T}_${WL}_$tag.json 2> gpurun_out/${T}_${WL}_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${T}_${WL}_$tag.json"))
    print("$WL $tag: %.2f M env-steps/s  %.3f ms  frac %.3f  launches %s %s" % (d["value"]/1e6, d["ms_per_step"], d["roofline"]["frac"], d["config"]["kernels_per_step"], d["config"]["rules_kernels"]))
except Exception as ex:
    print("$WL $tag: FAILED", ex)
PY
}
for WL in cfg2_rollout16 diag_wrush16 diag_lrush16 cfg1_selfplay16 cfg3_selfplay8 cfg4_partial32 cfg5_bloodbath64; do
  one auto MRTS_PIPELINE=1; one bulk0 MRTS_PIPELINE=1 MRTS_BULK_OUT=0; one bulk2 MRTS_PIPELINE=1 MRTS_BULK_OUT=2; one bulk3 MRTS_PIPELINE=1 MRTS_BULK_OUT=3
done 2>&1 | tee -a gpurun_out/${T}_summary.txt
MRTS_PIPELINE=1 timeout 300 python tools/game_timing.py diag_wrush16 300 > gpurun_out/${T}_timing_wrush.txt 2>&1
tail -n 34 gpurun_out/${T}_timing_wrush.txt
