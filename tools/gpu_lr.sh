set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="--steps 50 --warmup 5 --cpu-budget 0 --e2e-envs 64"
for w in cfg2_rollout16 diag_wrush16 diag_lrush16 diag_mix3_16 cfg1_selfplay16 cfg3_selfplay8; do python bench.py --workload $w $B > gpurun_out/r2l_$w.json 2> gpurun_out/r2l_$w.err; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2l_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, "%.2fM"%(d["value"]/1e6), "%.3f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"])
    except Exception as e: print(f, "ERR", e)
PY
