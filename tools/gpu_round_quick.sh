# Trimmed round measurement (about 9 GPU-minutes): GPU tests, both arms of the default bench line, the other
# workloads at 100 steps, the ncu launch list and one full capture of the default workload.
# Usage: bash tools/gpu_round_quick.sh <tag>
set -x
cd $GRAFT_REPO_ROOT
T=${1:-r01i}
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
timeout 300 python bench.py > gpurun_out/${T}_bench_default.json 2> gpurun_out/${T}_bench_default.err
timeout 300 python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
B="--steps 100 --warmup 10 --cpu-budget 0"
for w in cfg1_selfplay16 cfg3_selfplay8 cfg4_partial32 cfg5_bloodbath64 diag_passive16 diag_wrush16; do timeout 200 python bench.py --workload $w $B > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err; done
timeout 300 python bench.py --dtype uint8 $B > gpurun_out/${T}_bench_cfg2_uint8.json 2> gpurun_out/${T}_bench_cfg2_uint8.err
timeout 300 python bench.py --workload cfg3_selfplay8 --dtype uint8 $B > gpurun_out/${T}_bench_cfg3_uint8.json 2> gpurun_out/${T}_bench_cfg3_uint8.err
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:mrts_ --launch-skip 660 -c 120 --csv --log-file gpurun_out/${T}_launches_cfg2.csv python bench.py --steps 40 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/${T}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mrts_pass_ --launch-skip 330 --launch-count 1 -o gpurun_out/${T}_cfg2_full -f python bench.py --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/${T}_ncu_full.log 2>&1
