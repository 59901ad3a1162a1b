#!/bin/bash
# Sweep of the line-FFT CTA shapes / radix lists (environment knobs of es_fft.cu) on one workload.
# usage: profiles/sweep_fft_cfg.sh C2 grid
WL=${1:-C2}; DIR=${2:-grid}
for cfg in "0,0,2" "0,0,1" "0,256,1" "0,512,1" "0,0,0" "0,256,2" "256,0,2" "64,0,2"; do
  for rad in "" "22,20,14" "20,14,22" "28,20,11" "16,11,7,5"; do
    echo "== cfg=$cfg radices=$rad"
    SKA_SDP_FUNC_ES_FFT_CFG=$cfg SKA_SDP_FUNC_ES_FFT_RADICES=$rad timeout 120 python bench.py --workload $WL --direction $DIR --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],2), {k:round(v['ms'],2) for k,v in d['stages'].items()})"
    [ "$cfg" != "0,0,2" ] && break
  done
done
