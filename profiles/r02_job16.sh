#!/bin/bash
# GPU-box job 16 (round 2): fp64 scatter with per-record ES weights evaluated
# once per call: parity suite, C3 timings with and without.
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 ) > $O/pytest_gpu16.log 2>&1
tail -12 $O/pytest_gpu16.log
timeout 400 python profiles/run_path.py C3 grid 3 > $O/ab16_C3_grid_kw.txt 2>&1
echo "C3 grid kweights: $(tail -2 $O/ab16_C3_grid_kw.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
SKA_SDP_FUNC_ES_NO_KWEIGHTS=1 timeout 400 python profiles/run_path.py C3 grid 3 > $O/ab16_C3_grid_nokw.txt 2>&1
echo "C3 grid in-kernel: $(tail -2 $O/ab16_C3_grid_nokw.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
timeout 400 python profiles/run_path.py C1 grid 3 > $O/ab16_C1_grid.txt 2>&1
echo "C1 grid: $(tail -2 $O/ab16_C1_grid.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
