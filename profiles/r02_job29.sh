#!/bin/bash
# GPU-box job 29 (round 2): Hogbom CLEAN with grouped PSF loads.
O=gpurun_out
mkdir -p $O
timeout 400 python -m pytest tests/test_hogbom_clean.py -m gpu -q -x > $O/pytest_clean29.log 2>&1; tail -2 $O/pytest_clean29.log
for cap in 148; do
  echo "max CTAs $cap"
  SKA_SDP_FUNC_CLEAN_MAX_CTAS=$cap timeout 400 python profiles/run_clean.py --n 256 --cycles 2000 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
  SKA_SDP_FUNC_CLEAN_MAX_CTAS=$cap timeout 400 python profiles/run_clean.py --n 1024 --cycles 1000 --single 1 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
  SKA_SDP_FUNC_CLEAN_MAX_CTAS=$cap timeout 400 python profiles/run_clean.py --n 2048 --cycles 500 --single 0 --no-ref --no-oracle --steps 3 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
done
