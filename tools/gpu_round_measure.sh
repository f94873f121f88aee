# Round measurement: GPU tests, the default bench line, every workload, uint8 mode, the reference arm,
# the ncu launch list and one full ncu capture of the default workload.  Usage: bash tools/gpu_round_measure.sh <tag>
set -x
cd $GRAFT_REPO_ROOT
T=${1:-r01}
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
python bench.py > gpurun_out/${T}_bench_default.json 2> gpurun_out/${T}_bench_default.err
python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
B="--steps 100 --warmup 10 --cpu-budget 0"
for w in cfg1_selfplay16 cfg3_selfplay8 cfg4_partial32 cfg5_bloodbath64 diag_passive16 diag_wrush16; do python bench.py --workload $w $B > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err; done
python bench.py --dtype uint8 $B > gpurun_out/${T}_bench_cfg2_uint8.json 2> gpurun_out/${T}_bench_cfg2_uint8.err
python bench.py --workload cfg3_selfplay8 --dtype uint8 $B > gpurun_out/${T}_bench_cfg3_uint8.json 2> gpurun_out/${T}_bench_cfg3_uint8.err
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:mrts_ --launch-skip 660 -c 120 --csv --log-file gpurun_out/${T}_launches_cfg2.csv python bench.py --steps 40 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/${T}_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mrts_pass_ --launch-skip 330 --launch-count 1 -o gpurun_out/${T}_cfg2_full -f python bench.py --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/${T}_ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mrts_pass_ --launch-skip 330 --launch-count 1 -o gpurun_out/${T}_cfg1_full -f python bench.py --workload cfg1_selfplay16 --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/${T}_ncu_full1.log 2>&1
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${T}_bench_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, "%.2fM"%(d["value"]/1e6), "%.3f ms"%d["ms_per_step"], "frac %.3f"%d.get("roofline",{}).get("frac",0), "e2e %.0f"%d["e2e"]["value"], d.get("cpu_baseline",{}).get("value"))
    except Exception as e: print(f, "ERR", e)
PY
