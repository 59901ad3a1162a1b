#!/bin/bash
# GPU-box job 15 (round 2): the allocation-free Briggs entry point: its test
# alone in a fresh process, then the whole GPU suite (it runs last).
O=gpurun_out
mkdir -p $O
( time timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) > $O/pytest_no_alloc_alone.log 2>&1
tail -6 $O/pytest_no_alloc_alone.log
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 ) > $O/pytest_gpu15.log 2>&1
tail -12 $O/pytest_gpu15.log
