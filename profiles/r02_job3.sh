#!/bin/bash
# GPU-box job 3 (round 2): full parity suite on the per-(radix, kind) pass
# functions; A/B against the inlined build and radix-subset builds.
O=gpurun_out
mkdir -p $O
P=ska-sdp-func_b200
( time timeout 1200 python -m pytest tests -m gpu -q --durations=8 ) > $O/pytest_gpu3.log 2>&1
tail -12 $O/pytest_gpu3.log
for v in default inl c2set; do
  for d in grid degrid; do
    if [ $v = default ]; then L=$P/lib; else L=$P/lib_$v; fi
    SKA_SDP_FUNC_LIB_DIR=$L timeout 300 python profiles/run_path.py C2 $d 4 > $O/ab3_c2_${d}_$v.txt 2>&1
    echo "C2 $d $v: $(tail -2 $O/ab3_c2_${d}_$v.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
  done
done
for v in default c4set inl; do
  if [ $v = default ]; then L=$P/lib; else L=$P/lib_$v; fi
  SKA_SDP_FUNC_LIB_DIR=$L timeout 400 python profiles/run_path.py C4 grid 2 > $O/ab3_c4_grid_$v.txt 2>&1
  echo "C4 grid $v: $(tail -2 $O/ab3_c4_grid_$v.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
  SKA_SDP_FUNC_LIB_DIR=$L SKA_SDP_FUNC_ES_FFT_RADICES=16,14,11,10 SKA_SDP_FUNC_ES_FFT_CFG=1024,1024,0,0 timeout 400 python profiles/run_path.py C4 grid 2 > $O/ab3_c4_grid_wide_$v.txt 2>&1
  echo "C4 grid wide $v: $(tail -2 $O/ab3_c4_grid_wide_$v.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
done
SKA_SDP_FUNC_LIB_DIR=$P/lib timeout 400 python profiles/run_path.py C4 degrid 2 > $O/ab3_c4_degrid_default.txt 2>&1
echo "C4 degrid default: $(tail -2 $O/ab3_c4_degrid_default.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
