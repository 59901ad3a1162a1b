#!/bin/bash
# GPU-box job 23 (round 2): point-source DFT predict: parity (oracle, reference
# CUDA build), timings, launch list.
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_dft.py -m gpu -q -s > $O/pytest_dft23.log 2>&1; tail -5 $O/pytest_dft23.log
for p in 1 4; do
  timeout 900 python profiles/run_dft.py --pols $p > $O/r02_dft_p$p.json 2> $O/r02_dft_p$p.err; tail -c 1500 $O/r02_dft_p$p.json; tail -3 $O/r02_dft_p$p.err
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file $O/r02_launches_dft.csv python profiles/run_dft.py --pols 4 --steps 1 --no-ref > $O/ncu_dft.log 2>&1; tail -2 $O/ncu_dft.log | cut -c1-300
