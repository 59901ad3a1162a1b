#!/bin/bash
# GPU-box job 34 (round 2): final state: whole GPU suite, smoke, default bench.
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu34.log 2>&1; tail -3 $O/pytest_gpu34.log
timeout 300 python __graft_entry__.py smoke > $O/smoke34.log 2>&1; tail -1 $O/smoke34.log | cut -c1-200
timeout 900 python bench.py > $O/r02_bench_C4_n1_final.json 2> $O/bench34.err; tail -c 400 $O/r02_bench_C4_n1_final.json; tail -2 $O/bench34.err
