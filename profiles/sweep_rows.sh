#!/bin/bash
# Sweep of the row-kernel CTA shape: SKA_SDP_FUNC_ES_FFT_CFG="rows_threads,cols_threads,cols_lshift,rows_lshift"
WL=${1:-C2}
for cfg in "0,0,2,0" "0,0,2,1" "0,0,2,2" "128,0,2,1" "512,0,2,2" "256,0,2,2"; do
  for dir in grid degrid; do
    echo "== cfg=$cfg $dir"
    SKA_SDP_FUNC_ES_FFT_CFG=$cfg timeout 120 python bench.py --workload $WL --direction $dir --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python profiles/summ.py /dev/stdin
  done
done
