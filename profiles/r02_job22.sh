#!/bin/bash
# GPU-box job 22 (round 2): tiled functions after the per-channel table and
# the staged emit: parity, timings against the reference CUDA build, launch list.
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_tiled_functions.py tests/test_weighting.py tests/test_zz_no_alloc.py -m gpu -q -s > $O/pytest_tiled22.log 2>&1; tail -5 $O/pytest_tiled22.log
for s in 0 1; do
  timeout 600 python profiles/run_tiled.py --single $s > $O/r02_tiled_s$s.json 2> $O/r02_tiled_s$s.err; tail -c 1200 $O/r02_tiled_s$s.json; tail -3 $O/r02_tiled_s$s.err
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_tiled.csv python profiles/run_tiled.py --single 0 --steps 2 --no-ref > $O/ncu_tiled.log 2>&1; tail -2 $O/ncu_tiled.log | cut -c1-300
