#!/bin/bash
# GPU-box job 26 (round 2): Hogbom CLEAN: parity, timings against the reference
# CUDA build.
O=gpurun_out
mkdir -p $O
timeout 400 python -m pytest tests/test_hogbom_clean.py -m gpu -q -s -x > $O/pytest_clean26.log 2>&1; tail -6 $O/pytest_clean26.log
timeout 400 python profiles/run_clean.py --n 256 --cycles 2000 > $O/r02_clean_256_f64.json 2> $O/r02_clean_256_f64.err; tail -c 900 $O/r02_clean_256_f64.json; tail -3 $O/r02_clean_256_f64.err
timeout 400 python profiles/run_clean.py --n 1024 --cycles 1000 --single 1 > $O/r02_clean_1024_f32.json 2> $O/r02_clean_1024_f32.err; tail -c 900 $O/r02_clean_1024_f32.json; tail -3 $O/r02_clean_1024_f32.err
