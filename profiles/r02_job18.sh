#!/bin/bash
# GPU-box job 18 (round 2): fp64 scatter variants: quick parity + C3 timings.
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_contract.py -m gpu -x -q -k "fp64 or golden or split_heavy or back_to_back or c3 or determin" > $O/pytest_gpu18.log 2>&1; tail -3 $O/pytest_gpu18.log
timeout 400 python profiles/run_path.py C3 grid 3 > $O/ab18_C3_grid_kw.txt 2>&1
echo "C3 grid kweights: $(tail -2 $O/ab18_C3_grid_kw.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
