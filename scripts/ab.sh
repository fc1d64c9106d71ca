#!/bin/bash
# A/B driver for tuning variants on the GPU box: scripts/ab.sh <tag> <workloads...> -- <variant names (in lib/variants) or "base">
tag=$1; shift
wl=(); while [ "$1" != "--" ]; do wl+=("$1"); shift; done; shift
for v in "$@"; do
  for w in "${wl[@]}"; do
    if [ "$v" = base ]; then unset TPT_B200_LIB; else export TPT_B200_LIB=$PWD/tinypathtracer_b200/lib/variants/$v.so; fi
    timeout 300 python bench.py --workload $w --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_${v}_$w.log 2>&1
  done
done
