#!/bin/bash
# A/B over environment switches: scripts/ab_env.sh <tag> <workloads...> -- <NAME=VAL | base> ...
tag=$1; shift
wl=(); while [ "$1" != "--" ]; do wl+=("$1"); shift; done; shift
for v in "$@"; do
  for w in "${wl[@]}"; do
    if [ "$v" = base ]; then env python bench.py --workload $w --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_base_$w.log 2>&1
    else env "$v" python bench.py --workload $w --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_${v//=/_}_$w.log 2>&1; fi
  done
done
