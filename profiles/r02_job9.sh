#!/bin/bash
# GPU-box job 9 (round 2): cluster column kernels with batched loads, both
# directions: parity, timings for cluster sizes 2 / 4 / 0.
O=gpurun_out
mkdir -p $O
( time timeout 1200 python -m pytest tests/test_gpu_timed_shapes.py tests/test_gpu_parity.py -m gpu -q --durations=5 ) > $O/pytest_gpu9.log 2>&1
tail -7 $O/pytest_gpu9.log
for cs in 2 4 0; do
  for d in grid degrid; do
    SKA_SDP_FUNC_ES_FFT_CLUSTER=$cs timeout 400 python profiles/run_path.py C4 $d 2 > $O/ab9_C4_${d}_cs$cs.txt 2>&1
    echo "C4 $d cluster $cs: $(tail -2 $O/ab9_C4_${d}_cs$cs.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
  done
done
for cs in 2 0; do
  for d in grid degrid; do
    SKA_SDP_FUNC_ES_FFT_CLUSTER=$cs timeout 400 python profiles/run_path.py C3 $d 3 > $O/ab9_C3_${d}_cs$cs.txt 2>&1
    echo "C3 $d cluster $cs: $(tail -2 $O/ab9_C3_${d}_cs$cs.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
  done
done
SKA_SDP_FUNC_ES_FFT_CLUSTER=4 timeout 600 python -m pytest tests/test_gpu_timed_shapes.py -m gpu -q -k "c4 or c3" > $O/pytest_gpu9_cs4.log 2>&1; tail -3 $O/pytest_gpu9_cs4.log
