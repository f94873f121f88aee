#!/bin/bash
# bulk writers in the direct pass + the new expansion kernel: parity on both pipelines, then A/B
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_trajectory_parity.py tests/test_gpu_api.py tests/test_known_answers.py -m gpu -x -q > gpurun_out/r02d_pytest.log 2>&1; echo "pytest default rc=$?" | tee gpurun_out/r02d_summary.txt
MRTS_PIPELINE=2 MRTS_CHUNKS=2 timeout 900 python -m pytest tests/test_trajectory_parity.py tests/test_gpu_api.py -m gpu -x -q -k "cfg or partial8 or modes or lightrush16 or cycle or staged or full_size" > gpurun_out/r02d_pytest_split.log 2>&1; echo "pytest split rc=$?" | tee -a gpurun_out/r02d_summary.txt
timeout 1500 python -m pytest tests/test_parity_matrix.py tests/test_reference_consumers.py -m gpu -x -q > gpurun_out/r02d_pytest_matrix.log 2>&1; echo "pytest matrix rc=$?" | tee -a gpurun_out/r02d_summary.txt
tail -4 gpurun_out/r02d_pytest.log gpurun_out/r02d_pytest_split.log gpurun_out/r02d_pytest_matrix.log
B="python bench.py --steps 40 --warmup 5 --cpu-budget 0 --e2e-envs 0"
one() { tag=$1; shift
  env "$@" timeout 300 $B --workload $WL > gpurun_out/r02d_${WL}_$tag.json 2> gpurun_out/r02d_${WL}_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02d_${WL}_$tag.json"))
    print("$WL $tag: %.2f M env-steps/s  %.3f ms  frac %.3f  launches %s %s" % (d["value"]/1e6, d["ms_per_step"], d["roofline"]["frac"], d["config"]["kernels_per_step"], d["config"]["rules_kernels"]))
except Exception as ex:
    print("$WL $tag: FAILED", ex)
PY
}
for WL in cfg2_rollout16 diag_wrush16 cfg1_selfplay16 cfg3_selfplay8 cfg4_partial32 cfg5_bloodbath64; do
  one direct_plain MRTS_PIPELINE=1 MRTS_BULK_MASK=0; one direct_bulk MRTS_PIPELINE=1; one direct_bulk_g2 MRTS_PIPELINE=1 MRTS_GPB_LIGHT=2 MRTS_GPB_HEAVY=2
  one split_k1 MRTS_PIPELINE=2 MRTS_CHUNKS=1; one split_k2 MRTS_PIPELINE=2 MRTS_CHUNKS=2
done 2>&1 | tee -a gpurun_out/r02d_summary.txt
MRTS_PIPELINE=1 timeout 300 python tools/game_timing.py diag_wrush16 300 > gpurun_out/r02d_timing_wrush.txt 2>&1
tail -32 gpurun_out/r02d_timing_wrush.txt
N="python bench.py --steps 3 --warmup 2 --prewarm 200 --cpu-budget 0 --e2e-envs 0"
MRTS_PIPELINE=1 timeout 300 $N --workload cfg3_selfplay8 > gpurun_out/r02d_plain3.log 2>&1 && \
MRTS_PIPELINE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:mrts_rules_warp -s 205 -c 1 -f -o gpurun_out/r02d_cfg3_direct $N --workload cfg3_selfplay8 > gpurun_out/r02d_ncu3.log 2>&1
MRTS_PIPELINE=1 timeout 300 $N --workload cfg2_rollout16 > gpurun_out/r02d_plain2.log 2>&1 && \
MRTS_PIPELINE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:mrts_rules_warp -s 410 -c 2 -f -o gpurun_out/r02d_cfg2_direct $N --workload cfg2_rollout16 > gpurun_out/r02d_ncu2.log 2>&1
