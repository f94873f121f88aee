set -x
cd $GRAFT_REPO_ROOT
ncu --set full --clock-control none --import-source on -k regex:mrts_pass --launch-skip 330 --launch-count 1 -o gpurun_out/r1e_wrush_full -f python bench.py --workload diag_wrush16 --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/r1e_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mrts_pass --launch-skip 330 --launch-count 1 -o gpurun_out/r1e_passive_full -f python bench.py --workload diag_passive16 --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/r1e_ncu2.log 2>&1
