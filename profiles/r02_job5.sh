#!/bin/bash
# GPU-box job 5 (round 2): kernel sets for fixed stage lists: parity suite,
# timings against the generic set, ncu captures of the C2 / C4 column kernels.
O=gpurun_out
mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q --durations=5 ) > $O/pytest_gpu5.log 2>&1
tail -9 $O/pytest_gpu5.log
for w in C2 C4 C5; do
  for d in grid degrid; do
    timeout 400 python profiles/run_path.py $w $d 3 > $O/ab5_${w}_${d}_seq.txt 2>&1
    echo "$w $d seq: $(tail -2 $O/ab5_${w}_${d}_seq.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
  done
done
SKA_SDP_FUNC_ES_FFT_GENERIC=1 timeout 400 python profiles/run_path.py C5 grid 3 > $O/ab5_C5_grid_generic.txt 2>&1
echo "C5 grid generic: $(tail -2 $O/ab5_C5_grid_generic.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
timeout 400 python profiles/run_path.py C3 degrid 3 > $O/ab5_C3_degrid_seq.txt 2>&1
echo "C3 degrid seq: $(tail -2 $O/ab5_C3_degrid_seq.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
timeout 400 python profiles/run_path.py C3 grid 3 > $O/ab5_C3_grid_seq.txt 2>&1
echo "C3 grid seq: $(tail -2 $O/ab5_C3_grid_seq.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:es_fft_grid -s 6 -c 2 -o $O/r02_ncu_c2_grid_b -f python profiles/run_path.py C2 grid 3 > $O/ncu_c2.log 2>&1; tail -2 $O/ncu_c2.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:es_fft_grid -s 6 -c 2 -o $O/r02_ncu_c4_grid_b -f python profiles/run_path.py C4 grid 1 > $O/ncu_c4.log 2>&1; tail -2 $O/ncu_c4.log
