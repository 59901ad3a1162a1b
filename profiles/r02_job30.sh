#!/bin/bash
# GPU-box job 30 (round 2): major-cycle integration test, then the whole GPU
# suite and the smoke entry.
O=gpurun_out
mkdir -p $O
timeout 400 python -m pytest tests/test_major_cycle.py -m gpu -q -s -x > $O/pytest_major30.log 2>&1; tail -12 $O/pytest_major30.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu30.log 2>&1; tail -4 $O/pytest_gpu30.log
timeout 300 python __graft_entry__.py smoke > $O/smoke30.log 2>&1; tail -2 $O/smoke30.log | cut -c1-300
