# Final bench lines of the round: default workload (both arms), plus the N-GPU line when more than one GPU is visible.
# Usage: bash tools/gpu_final_lines.sh <tag>
set -x
cd $GRAFT_REPO_ROOT
T=${1:-r01h}
N=$(nvidia-smi -L | wc -l)
python bench.py > gpurun_out/${T}_bench_default.json 2> gpurun_out/${T}_bench_default.err
python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
python bench.py --workload cfg3_selfplay8 --steps 100 --warmup 10 --cpu-budget 0 > gpurun_out/${T}_bench_cfg3.json 2> gpurun_out/${T}_bench_cfg3.err
if [ "$N" -gt 1 ]; then
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err
fi
tail -c 600 gpurun_out/${T}_bench_default.json
