#!/bin/bash
# GPU-box job 7 (round 2): compile-time line plans + 640-thread C4 set:
# parity suite, timings; then the sanitizer runs (job 6).
O=gpurun_out
mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q --durations=5 ) > $O/pytest_gpu7.log 2>&1
tail -9 $O/pytest_gpu7.log
for w in C2 C4 C5; do
  for d in grid degrid; do
    timeout 400 python profiles/run_path.py $w $d 3 > $O/ab7_${w}_${d}.txt 2>&1
    echo "$w $d: $(tail -2 $O/ab7_${w}_${d}.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
  done
done
timeout 400 python profiles/run_path.py C3 degrid 3 > $O/ab7_C3_degrid.txt 2>&1
echo "C3 degrid: $(tail -2 $O/ab7_C3_degrid.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
bash profiles/r02_job6.sh
