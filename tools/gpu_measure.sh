#!/bin/bash
# One GPU-box script for every measurement of a round:  tools/gpu_measure.sh TAG STAGE [STAGE ...]
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/gpu_measure.sh r02g noalloc pytest ab'
# Everything lands in gpurun_out/TAG_*; TAG_summary.txt is the digest.  Stages:
#   noalloc   tests/test_zz_no_alloc.py alone in a fresh process
#   pytest    the whole GPU suite (-m gpu -x)
#   smoke     __graft_entry__.smoke()
#   ab        device-resident throughput of $WORKLOADS (default: the five BASELINE configs + diag_wrush16), $ENVV applied
#   bench     the default bench.py line (cpu_baseline + e2e included) for $WORKLOADS -> TAG_bench.jsonl
#   launches  ncu launch list (gpu__time_duration) of $NCU_WL (default cfg2_rollout16)
#   full      ncu --set full of the dominant kernel of $NCU_WL -> TAG_full.ncu-rep (+ traffic json via tools/ncu_traffic.py)
#   scale     NGPU=N torchrun bench lines of $WORKLOADS (run under gpurun --gpus N)
#   src       ncu source counters of the heavy rules kernel on diag_wrush16
# ncu stages run the same command without ncu first and only go on when it exits 0.
set -u
T=${1:?tag}; shift
mkdir -p gpurun_out
S=gpurun_out/${T}_summary.txt
WORKLOADS=${WORKLOADS:-"cfg2_rollout16 diag_wrush16 cfg1_selfplay16 cfg3_selfplay8 cfg4_partial32 cfg5_bloodbath64"}
NCU_WL=${NCU_WL:-cfg2_rollout16}
ENVV=${ENVV:-}
EXTRA=${EXTRA:-}
say() { echo "$@" | tee -a $S; }

digest() {  # json file, label
python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r, c = d.get("roofline") or {}, d.get("config", {})
    e = d.get("e2e") or {}
    print("%s: %.2f M %s  %.3f ms/step  frac %.3f  e2e %s  launches %s" % (sys.argv[2], d["value"] / 1e6, d["unit"], d["ms_per_step"], r.get("frac", 0), ("%.2f M" % (e["value"] / 1e6)) if e.get("value") else "-", d.get("gpu_launches")))
except Exception as ex:
    print(sys.argv[2] + ": FAILED", ex)
PY
}

for stage in "$@"; do
case $stage in
noalloc)
  timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/${T}_noalloc.log 2>&1; say "noalloc alone rc=$? $(tail -n 1 gpurun_out/${T}_noalloc.log)";;
pytest)
  timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; say "pytest -m gpu rc=$? $(tail -n 1 gpurun_out/${T}_pytest.log)";;
smoke)
  timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; say "smoke rc=$? $(tail -n 1 gpurun_out/${T}_smoke.log)";;
ab)
  for WL in $WORKLOADS; do
    env $ENVV timeout 400 python bench.py --steps 40 --warmup 5 --cpu-budget 0 --e2e-envs 0 --workload $WL $EXTRA > gpurun_out/${T}_ab_$WL.json 2> gpurun_out/${T}_ab_$WL.err
    say "$(digest gpurun_out/${T}_ab_$WL.json "$WL [$ENVV $EXTRA]")"
  done;;
bench)
  : > gpurun_out/${T}_bench.jsonl
  for WL in $WORKLOADS; do
    env $ENVV timeout 900 python bench.py --workload $WL $EXTRA > gpurun_out/${T}_bench_$WL.json 2> gpurun_out/${T}_bench_$WL.err
    tail -n 1 gpurun_out/${T}_bench_$WL.json >> gpurun_out/${T}_bench.jsonl
    say "$(digest gpurun_out/${T}_bench_$WL.json "bench $WL [$EXTRA]")"
  done;;
launches)
  B="python bench.py --steps 4 --warmup 3 --prewarm 100 --cpu-budget 0 --e2e-envs 0 --workload $NCU_WL $EXTRA"
  timeout 400 $B > gpurun_out/${T}_launches_plain.log 2>&1 && \
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_$NCU_WL.csv $B > gpurun_out/${T}_launches_ncu.log 2>&1
  say "launches rc=$? $(wc -l < gpurun_out/${T}_launches_$NCU_WL.csv 2>/dev/null) csv lines";;
full)
  B="python bench.py --steps 4 --warmup 3 --prewarm 300 --cpu-budget 0 --e2e-envs 0 --workload $NCU_WL $EXTRA"
  timeout 400 $B > gpurun_out/${T}_full_plain.log 2>&1 && \
  timeout 1800 ncu --set full --clock-control none --import-source on -k regex:mrts_ -s ${NCU_SKIP:-906} -c ${NCU_COUNT:-6} -f -o gpurun_out/${T}_full_$NCU_WL $B > gpurun_out/${T}_full_ncu.log 2>&1
  say "full rc=$? $(ls -la gpurun_out/${T}_full_$NCU_WL.ncu-rep 2>&1 | awk '{print $5}') bytes";;
src)
  B="python bench.py --steps 12 --warmup 2 --prewarm 200 --cpu-budget 0 --e2e-envs 0 --workload diag_wrush16"
  timeout 400 $B > gpurun_out/${T}_src_plain.log 2>&1 && \
  timeout 1500 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --clock-control none --import-source on -k regex:mrts_rules_warp -s 205 -c 12 -f -o gpurun_out/${T}_src_wrush $B > gpurun_out/${T}_src_ncu.log 2>&1
  say "src rc=$?";;
scale)  # under `gpurun --gpus N`: NGPU=N; one torchrun per workload, e2e arm included, CPU legs skipped
  : ${NGPU:?NGPU}
  for WL in $WORKLOADS; do
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $NGPU --workload $WL --steps 100 --warmup 10 --cpu-budget 0 $EXTRA > gpurun_out/${T}_scale${NGPU}_$WL.json 2> gpurun_out/${T}_scale${NGPU}_$WL.err
    tail -n 1 gpurun_out/${T}_scale${NGPU}_$WL.json >> gpurun_out/${T}_scale.jsonl
    say "$(digest gpurun_out/${T}_scale${NGPU}_$WL.json "N=$NGPU $WL [$EXTRA]")"
  done;;
*) say "unknown stage $stage";;
esac
done
cat $S
