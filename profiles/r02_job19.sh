#!/bin/bash
# GPU-box job 19 (round 2): scatter work-list order (plane-major / tile-major).
O=gpurun_out
mkdir -p $O
for o in 1 2; do
  SKA_SDP_FUNC_ES_UNIT_ORDER=$o timeout 400 python profiles/run_path.py C3 grid 3 > $O/ab19_C3_grid_o$o.txt 2>&1
  echo "C3 grid order $o: $(tail -2 $O/ab19_C3_grid_o$o.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
done
for o in 1 2; do
  SKA_SDP_FUNC_ES_UNIT_ORDER=$o timeout 400 python profiles/run_path.py C2 grid 4 > $O/ab19_C2_grid_o$o.txt 2>&1
  echo "C2 grid order $o: $(tail -2 $O/ab19_C2_grid_o$o.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
done
