#!/bin/bash
# compute-sanitizer over the kernels with cross-CTA protocols (race detection): the CTA-pair fp16-plane GEMM
# (cta_group::2, remote mbarrier arrives, TMA multicast commit), the 8-CTA-cluster LSTM recurrence (cluster barriers, partial
# sums exchanged through L2) and the plane epilogues, on small cases.  ONE tool per gpurun call (B200_PROFILING.md):
#   gpurun -- bash scripts/sanitize.sh racecheck     (or memcheck / synccheck)
tool=${1:-racecheck}
compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_case.py > gpurun_out/sanitize_$tool.log 2>&1
echo "compute-sanitizer $tool rc=$?" >> gpurun_out/sanitize_$tool.log
tail -15 gpurun_out/sanitize_$tool.log
