#!/bin/bash
# register-budget / games-per-block A/B in direct mode, per-game phase timing, ncu of the heavy and light rules kernels
mkdir -p gpurun_out
B="python bench.py --steps 40 --warmup 5 --cpu-budget 0 --e2e-envs 0"
one() { # tag, env..., -- args
  tag=$1; shift
  env "$@" MRTS_PIPELINE=1 timeout 300 $B --workload $WL > gpurun_out/r02b_${WL}_$tag.json 2> gpurun_out/r02b_${WL}_$tag.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r02b_${WL}_$tag.json"))
    print("$WL $tag: %.2f M env-steps/s  %.3f ms  frac %.3f  %s" % (d["value"]/1e6, d["ms_per_step"], d["roofline"]["frac"], d["config"]["rules_kernels"]))
except Exception as ex:
    print("$WL $tag: FAILED", ex)
PY
}
for WL in diag_wrush16 cfg2_rollout16; do
  one h64 MRTS_REGS_HEAVY=64; one h96 MRTS_REGS_HEAVY=96; one h128 MRTS_REGS_HEAVY=128
  one h64g1 MRTS_REGS_HEAVY=64 MRTS_GPB_HEAVY=1; one h64g2 MRTS_REGS_HEAVY=64 MRTS_GPB_HEAVY=2; one h96g1 MRTS_REGS_HEAVY=96 MRTS_GPB_HEAVY=1; one h96g2 MRTS_REGS_HEAVY=96 MRTS_GPB_HEAVY=2
done
for WL in cfg1_selfplay16 cfg3_selfplay8 diag_passive16; do
  one l64 MRTS_REGS_LIGHT=64; one l96 MRTS_REGS_LIGHT=96; one l64g1 MRTS_REGS_LIGHT=64 MRTS_GPB_LIGHT=1; one l64g2 MRTS_REGS_LIGHT=64 MRTS_GPB_LIGHT=2
done
MRTS_PIPELINE=1 timeout 300 python tools/game_timing.py diag_wrush16 300 > gpurun_out/r02b_timing_wrush.txt 2>&1
MRTS_PIPELINE=1 timeout 300 python tools/game_timing.py diag_passive16 300 > gpurun_out/r02b_timing_passive.txt 2>&1
tail -20 gpurun_out/r02b_timing_wrush.txt
# ncu: heavy kernel, direct mode (fused) and rules only (split, 1 chunk)
N="python bench.py --steps 3 --warmup 2 --prewarm 200 --cpu-budget 0 --e2e-envs 0"
MRTS_PIPELINE=1 timeout 300 $N --workload diag_wrush16 > gpurun_out/r02b_plain1.log 2>&1 && \
MRTS_PIPELINE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:mrts_rules_warp -s 205 -c 1 -f -o gpurun_out/r02b_wrush_direct $N --workload diag_wrush16 > gpurun_out/r02b_ncu1.log 2>&1
MRTS_PIPELINE=2 MRTS_CHUNKS=1 timeout 300 $N --workload diag_wrush16 > gpurun_out/r02b_plain2.log 2>&1 && \
MRTS_PIPELINE=2 MRTS_CHUNKS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"mrts_rules_warp|mrts_expand" -s 408 -c 2 -f -o gpurun_out/r02b_wrush_split $N --workload diag_wrush16 > gpurun_out/r02b_ncu2.log 2>&1
MRTS_PIPELINE=1 timeout 300 $N --workload cfg1_selfplay16 > gpurun_out/r02b_plain3.log 2>&1 && \
MRTS_PIPELINE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:mrts_rules_warp -s 205 -c 1 -f -o gpurun_out/r02b_cfg1_direct $N --workload cfg1_selfplay16 > gpurun_out/r02b_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep
