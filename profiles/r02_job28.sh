#!/bin/bash
# GPU-box job 28 (round 2): Hogbom CLEAN: CTAs at the grid barrier.
O=gpurun_out
mkdir -p $O
for cap in 16 32 64 96 128 148; do
  echo "max CTAs $cap"
  for rep in 1 2; do
  SKA_SDP_FUNC_CLEAN_MAX_CTAS=$cap timeout 400 python profiles/run_clean.py --n 256 --cycles 2000 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
  done
  SKA_SDP_FUNC_CLEAN_MAX_CTAS=$cap timeout 400 python profiles/run_clean.py --n 1024 --cycles 1000 --single 1 --no-ref --no-oracle --steps 5 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
  SKA_SDP_FUNC_CLEAN_MAX_CTAS=$cap timeout 400 python profiles/run_clean.py --n 2048 --cycles 500 --single 0 --no-ref --no-oracle --steps 3 2>/dev/null | grep -o '"us_per_cycle": [0-9.]*'
done
