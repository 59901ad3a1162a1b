#!/bin/bash
# GPU-box job 14 (round 2): stage orders with fewer shared-memory bank conflicts
# (kernel sets 5-9), timed against the planner's default orders.
O=gpurun_out
mkdir -p $O
run() { # workload direction calls radices tag
  if [ -n "$4" ]; then export SKA_SDP_FUNC_ES_FFT_RADICES=$4; else unset SKA_SDP_FUNC_ES_FFT_RADICES; fi
  SKA_SDP_FUNC_ES_DEBUG=1 timeout 400 python profiles/run_path.py $1 $2 $3 > $O/ab14_$1_$2_$5.txt 2>&1
  echo "$1 $2 $5: $(grep -o "kernel set [0-9]*" $O/ab14_$1_$2_$5.txt | head -1) $(tail -2 $O/ab14_$1_$2_$5.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
}
for d in grid degrid; do
  run C2 $d 4 "" default
  run C2 $d 4 28,11,20 r28_11_20
  run C2 $d 4 20,11,28 r20_11_28
done
for d in grid degrid; do
  run C5 $d 2 "" default
  run C5 $d 2 24,15,32 r24_15_32
done
for d in grid degrid; do
  run C4 $d 2 "" default
  run C4 $d 2 14,11,10,16 r14_11_10_16
  run C4 $d 2 10,11,14,16 r10_11_14_16
done
