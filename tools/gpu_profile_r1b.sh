set -x
cd $GRAFT_REPO_ROOT
for w in diag_passive16 diag_wrush16; do python bench.py --workload $w --steps 50 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/r1b_bench_$w.json 2> gpurun_out/r1b_bench_$w.err; cat gpurun_out/r1b_bench_$w.json; done
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:mrts_pass --launch-skip 300 -c 60 --csv --log-file gpurun_out/r1b_launches_cfg2.csv python bench.py --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/r1b_ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mrts_pass --launch-skip 330 --launch-count 1 -o gpurun_out/r1b_cfg2_full -f python bench.py --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/r1b_ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mrts_pass --launch-skip 330 --launch-count 1 -o gpurun_out/r1b_cfg1_full -f python bench.py --workload cfg1_selfplay16 --steps 30 --warmup 5 --cpu-budget 0 --e2e-envs 64 > gpurun_out/r1b_ncu_full1.log 2>&1
ls -la gpurun_out
