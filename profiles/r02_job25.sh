#!/bin/bash
# GPU-box job 25 (round 2): ncu --set full of the DFT kernels (P = 1).
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:es_dft_kernel -c 16 -o $O/r02_ncu_dft_p1 -f python profiles/run_dft.py --pols 1 --steps 1 --no-ref > $O/ncu_dft_full.log 2>&1; tail -2 $O/ncu_dft_full.log | cut -c1-200
ncu -i $O/r02_ncu_dft_p1.ncu-rep --page raw --csv > $O/r02_ncu_dft_p1_raw.csv 2>/dev/null; wc -c $O/r02_ncu_dft_p1_raw.csv
