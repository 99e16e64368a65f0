#!/bin/bash
# one `ncu --set full` capture per hot kernel of the cfg3 step (steady state: skips the warm-up launches of that kernel)
# usage (on the GPU box): bash scripts/ncu_kernels.sh <tag> "<kernel regex> <skip>" ...
tag=$1; shift
for spec in "$@"; do
  set -- $spec
  k=$1; skip=$2
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip $skip --launch-count 1 \
      -f -o gpurun_out/${tag}_${k} python bench.py --steps 1 --warmup 3 --no-roofline --no-cpu-baseline > gpurun_out/${tag}_${k}.log 2>&1
  echo "$k rc=$?"
done
