#!/bin/bash
# GPU-box job 17 (round 2): ncu capture of the fp64 scatter on C3.
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:es_scatter2 -s 2 -c 2 -o $O/r02_ncu_c3_scatter -f python profiles/run_path.py C3 grid 2 > $O/ncu_c3s.log 2>&1; tail -3 $O/ncu_c3s.log
