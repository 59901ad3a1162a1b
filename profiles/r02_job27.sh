#!/bin/bash
# GPU-box job 27 (round 2): Hogbom CLEAN with incremental indexing; elements
# per thread (= number of CTAs at the grid barrier) 2 / 4 / 8 / 16.
O=gpurun_out
mkdir -p $O
timeout 400 python -m pytest tests/test_hogbom_clean.py -m gpu -q -x > $O/pytest_clean27.log 2>&1; tail -2 $O/pytest_clean27.log
for pt in 2 4 8 16; do
  echo "per thread $pt"
  SKA_SDP_FUNC_CLEAN_PER_THREAD=$pt timeout 400 python profiles/run_clean.py --n 256 --cycles 2000 --no-ref --no-oracle 2>/dev/null | grep -o '"ours": {[^}]*}'
  SKA_SDP_FUNC_CLEAN_PER_THREAD=$pt timeout 400 python profiles/run_clean.py --n 1024 --cycles 1000 --single 1 --no-ref --no-oracle 2>/dev/null | grep -o '"ours": {[^}]*}'
done
