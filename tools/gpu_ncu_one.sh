# usage: gpu_ncu_one.sh <workload> <tag>   -- one full ncu capture of the step kernel after the pre-warm
set -x
cd $GRAFT_REPO_ROOT
ncu --set full --clock-control none --import-source on -k regex:mrts_pass_ --launch-skip 330 --launch-count 1 -o gpurun_out/$2 -f python bench.py --workload $1 --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/$2.log 2>&1
