set -x
cd $GRAFT_REPO_ROOT
B="--steps 50 --warmup 5 --cpu-budget 0 --e2e-envs 64"
for m in 6 8; do for w in cfg2_rollout16 diag_wrush16 cfg1_selfplay16 cfg4_partial32; do MRTS_B200_LIB=$PWD/microrts_py_b200/csrc/variants/libmrts_b200_minb$m.so python bench.py --workload $w $B > gpurun_out/r2d_${w}_minb$m.json 2> gpurun_out/r2d_${w}_minb$m.err; done; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2d_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, "%.2fM"%(d["value"]/1e6), "%.3f ms"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"])
    except Exception as e: print(f, "ERR", e)
PY
