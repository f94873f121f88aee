# A/B of one environment variable over the main workloads: tools/gpu_ab.sh VAR "v1 v2 ..." [tag]
set -x
cd $GRAFT_REPO_ROOT
VAR=$1; VALS=$2; TAG=${3:-ab}
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="--steps 50 --warmup 5 --cpu-budget 0 --e2e-envs 64"
for v in $VALS; do
for w in cfg2_rollout16 diag_passive16 diag_wrush16 cfg1_selfplay16 cfg3_selfplay8 cfg4_partial32; do env $VAR=$v python bench.py --workload $w $B > gpurun_out/${TAG}_${v}_$w.json 2> gpurun_out/${TAG}_${v}_$w.err; done
done
python - $TAG <<'PY'
import json,glob,sys
for f in sorted(glob.glob("gpurun_out/%s_*.json"%sys.argv[1])):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, "%.2fM"%(d["value"]/1e6), "%.3f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"])
    except Exception as e: print(f, "ERR", e)
PY
