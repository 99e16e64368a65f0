#!/bin/bash
# one `ncu --set full` capture per hot kernel (steady state: --launch-skip passes the warm-up launches of that kernel); the step
# runs eagerly (B200IQ_GRAPH=0) so that every launch is a plain kernel launch.  Usage (GPU box): bash scripts/ncu_full.sh <tag>
tag=${1:-r02}
run() {  # name, workload, skip, kernel regex (demangled)
  local name=$1 wl=$2 skip=$3 rx=$4
  B200IQ_GRAPH=0 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$rx" --launch-skip $skip --launch-count 1 \
      -f -o gpurun_out/${tag}_${name} python bench.py --workload $wl --steps 1 --warmup 3 --no-roofline --no-cpu-baseline --no-secondary > gpurun_out/${tag}_${name}.log 2>&1
  echo "$name rc=$?"
}
B200IQ_GRAPH=0 python bench.py --steps 1 --warmup 3 --no-roofline --no-cpu-baseline --no-secondary > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
run gemm_fwd3_ffn transformer 40 'gemm_f16_kernel<.int.256, .bool.0, .bool.0, .int.3, .int.3, .int.2, .int.2>'
run gemm_dx transformer 100 'gemm_f16_kernel<.int.256, .bool.0, .bool.1, .int.1, .int.5, .int.0, .int.2>'
run gemm_dw transformer 150 'gemm_f16_kernel<.int.256, .bool.1, .bool.1, .int.1, .int.5, .int.0, .int.2>'
run attention_fwd transformer 20 'attention_fwd_tc_kernel<.int.2, .int.3>'
run attention_bwd transformer 20 'attention_bwd_tc_kernel<.int.2, .int.3>'
run ln_bwd transformer 50 'ln_bwd_warp_kernel<.int.4>'
if [ "$2" != "nolstm" ]; then
run lstm_seq_fwd lstm 5 'lstm_seq_fwd_kernel'
run lstm_seq_bwd lstm 4 'lstm_seq_bwd_kernel'
fi
ls -la gpurun_out/${tag}_*.ncu-rep
