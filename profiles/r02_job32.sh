#!/bin/bash
# GPU-box job 32 (round 2): Hogbom CLEAN: one-cluster path (DSMEM candidate
# exchange) against the grid path.
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_hogbom_clean.py tests/test_major_cycle.py -m gpu -q -x > $O/pytest_clean32.log 2>&1; tail -3 $O/pytest_clean32.log
SKA_SDP_FUNC_CLEAN_NO_CLUSTER=1 timeout 300 python -m pytest tests/test_hogbom_clean.py -m gpu -q -x > $O/pytest_clean32_grid.log 2>&1; tail -2 $O/pytest_clean32_grid.log
for nc in 0 1; do
  echo "no cluster $nc"
  for n in 128 256 512; do
  SKA_SDP_FUNC_CLEAN_NO_CLUSTER=$nc timeout 200 python profiles/run_clean.py --n $n --cycles 2000 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
  done
  SKA_SDP_FUNC_CLEAN_NO_CLUSTER=$nc timeout 200 python profiles/run_clean.py --n 512 --single 1 --cycles 2000 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
done
