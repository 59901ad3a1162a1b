#!/bin/bash
# GPU-box job 37 (round 2): CLEAN at 512^2 fp64: cluster path (4 PSF groups per
# thread) against the grid path, per-cycle cost from two cycle counts.
for g in 4 1; do
  echo "cluster groups $g"
  for cyc in 2000 6000; do
  SKA_SDP_FUNC_CLEAN_CLUSTER_GROUPS=$g timeout 100 python profiles/run_clean.py --n 512 --cycles $cyc --no-ref --no-oracle --steps 3 2>/dev/null | grep -o '"ms": [0-9.]*'
  done
done
