#!/bin/bash
# GPU-box job 11 (round 2): weighting kernels + full suite; degrid-row skip
# conditional on sparsity.
O=gpurun_out
mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q --durations=5 ) > $O/pytest_gpu11.log 2>&1
tail -12 $O/pytest_gpu11.log
for w in C2 C4 C5; do
  timeout 400 python profiles/run_path.py $w degrid 3 > $O/ab11_${w}_degrid.txt 2>&1
  echo "$w degrid: $(tail -2 $O/ab11_${w}_degrid.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
done
