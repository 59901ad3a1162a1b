#!/bin/bash
# GPU-box job 36 (round 2): ncu --set full with source of the CLEAN cycles kernel.
O=gpurun_out
mkdir -p $O
timeout 300 ncu --set full --clock-control none --import-source on -k regex:es_hogbom_cycles -c 1 -o $O/r02_ncu_clean_256 -f python profiles/run_clean.py --n 256 --cycles 2000 --steps 0 --no-ref --no-oracle > $O/ncu_clean.log 2>&1; tail -2 $O/ncu_clean.log | cut -c1-200
ncu -i $O/r02_ncu_clean_256.ncu-rep --page raw --csv > $O/r02_ncu_clean_256_raw.csv 2>/dev/null
ncu -i $O/r02_ncu_clean_256.ncu-rep --page source --csv > $O/r02_ncu_clean_256_source.csv 2>/dev/null; wc -c $O/r02_ncu_clean_256_*.csv
