#!/bin/bash
# first GPU pass of the round: parity suite on the default pipeline, the split pipeline forced on the small cases, A/B bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02a_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1; echo "pytest default rc=$?" | tee -a gpurun_out/r02a_summary.txt
MRTS_PIPELINE=2 MRTS_CHUNKS=2 timeout 900 python -m pytest tests/test_trajectory_parity.py tests/test_gpu_api.py -m gpu -x -q -k "cfg or partial8 or modes or lightrush16 or cycle or staged or full_size" > gpurun_out/r02a_pytest_split.log 2>&1; echo "pytest split rc=$?" | tee -a gpurun_out/r02a_summary.txt
tail -3 gpurun_out/r02a_pytest.log gpurun_out/r02a_pytest_split.log
tools/gpu_ab.sh r02a cfg2_rollout16 diag_wrush16 diag_passive16 cfg1_selfplay16 cfg3_selfplay8 2>&1 | tee -a gpurun_out/r02a_summary.txt
