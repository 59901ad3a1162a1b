#!/bin/bash
# GPU-box job 31 (round 2): Hogbom CLEAN: own grid barrier against
# cooperative_groups' grid.sync().
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_hogbom_clean.py tests/test_major_cycle.py -m gpu -q -x > $O/pytest_clean31.log 2>&1; tail -2 $O/pytest_clean31.log
for cg in 0 1; do
  echo "cg barrier $cg"
  SKA_SDP_FUNC_CLEAN_CG_BARRIER=$cg timeout 200 python profiles/run_clean.py --n 256 --cycles 2000 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
  SKA_SDP_FUNC_CLEAN_CG_BARRIER=$cg timeout 200 python profiles/run_clean.py --n 1024 --cycles 1000 --single 1 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
  SKA_SDP_FUNC_CLEAN_CG_BARRIER=$cg timeout 200 python profiles/run_clean.py --n 2048 --cycles 500 --single 0 --no-ref --no-oracle --steps 3 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
done
