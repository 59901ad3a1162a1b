#!/bin/bash
# GPU-box job 8 (round 2): cluster (DSMEM) column kernel for single-line CTAs:
# parity at the C4 / C3 shapes, timings for cluster sizes 0 / 2 / 4.
O=gpurun_out
mkdir -p $O
( time timeout 1200 python -m pytest tests/test_gpu_timed_shapes.py tests/test_gpu_parity.py -m gpu -q --durations=5 ) > $O/pytest_gpu8.log 2>&1
tail -7 $O/pytest_gpu8.log
for cs in 2 4 0; do
  SKA_SDP_FUNC_ES_FFT_CLUSTER=$cs timeout 400 python profiles/run_path.py C4 grid 2 > $O/ab8_C4_grid_cs$cs.txt 2>&1
  echo "C4 grid cluster $cs: $(tail -2 $O/ab8_C4_grid_cs$cs.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
done
for cs in 2 0; do
  SKA_SDP_FUNC_ES_FFT_CLUSTER=$cs timeout 400 python profiles/run_path.py C3 grid 3 > $O/ab8_C3_grid_cs$cs.txt 2>&1
  echo "C3 grid cluster $cs: $(tail -2 $O/ab8_C3_grid_cs$cs.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
done
SKA_SDP_FUNC_ES_FFT_CLUSTER=4 timeout 600 python -m pytest tests/test_gpu_timed_shapes.py -m gpu -q -k "c4 or c3" > $O/pytest_gpu8_cs4.log 2>&1; tail -3 $O/pytest_gpu8_cs4.log
