#!/bin/bash
# ncu --set full capture of the dominant kernel of each BASELINE config (tools/bench_configs.py), one timed frame each, after the same
# command has exited 0 without ncu.  Run under gpurun from the repo root; reports land in gpurun_out/.
#   -s: launches of that kernel before the first timed frame (instrumented counting launches + warm-up)
set -u
mkdir -p gpurun_out
cap() {   # key  kernel-regex  skip
  python tools/bench_configs.py --only "$1" --steps 2 --warmup 1 > "gpurun_out/r2_plain_$1.log" 2>&1 &&
  ncu --set full --clock-control none --import-source on -k "regex:$2" -s "$3" -c 1 -f -o "gpurun_out/r2_ncu_$1" \
      python tools/bench_configs.py --only "$1" --steps 2 --warmup 1 > "gpurun_out/r2_ncu_$1.log" 2>&1
  echo "$1 rc=$?"
}
for k in "$@"; do
  case "$k" in
    c1)   cap c1 rc1pass_kernel 2 ;;
    c3)   cap c3 shaded_march_kernel 3 ;;
    c3s)  cap c3s shaded_march_kernel 3 ;;
    c4)   cap c4 ebs_kernel 4 ;;
    vct)  cap vct shaded_march_kernel 4 ;;
    sbtm) cap sbtm sbtm_kernel 2 ;;
  esac
done
