#!/bin/bash
# GPU-box job 33 (round 2): Hogbom CLEAN: model update off the critical path;
# per-cycle cost from two cycle counts (fixed cost of a call separated).
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_hogbom_clean.py tests/test_major_cycle.py -m gpu -q -x > $O/pytest_clean33.log 2>&1; tail -2 $O/pytest_clean33.log
for nc in 0 1; do
  echo "no cluster $nc"
  for cyc in 2000 8000; do
  SKA_SDP_FUNC_CLEAN_NO_CLUSTER=$nc timeout 200 python profiles/run_clean.py --n 256 --cycles $cyc --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"ms": [0-9.]*'
  done
done
SKA_SDP_FUNC_CLEAN_NO_CLUSTER=1 timeout 200 python profiles/run_clean.py --n 1024 --single 1 --cycles 1000 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"ms": [0-9.]*'
SKA_SDP_FUNC_CLEAN_NO_CLUSTER=1 timeout 200 python profiles/run_clean.py --n 1024 --single 1 --cycles 4000 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"ms": [0-9.]*'
