# The default bench line (both arms) and its ncu launch list.  Usage: bash tools/gpu_default_line.sh <tag>
set -x
cd $GRAFT_REPO_ROOT
T=${1:-r01j}
( time timeout 400 python bench.py > gpurun_out/${T}_bench_default.json 2> gpurun_out/${T}_bench_default.err ) 2> gpurun_out/${T}_time.txt
timeout 200 python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
tail -c 300 gpurun_out/${T}_bench_default.json; cat gpurun_out/${T}_time.txt
