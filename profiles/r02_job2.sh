#!/bin/bash
# GPU-box job 2 (round 2): table-form w-screen parity, C4 parity diagnosis,
# A/B of the line-FFT build variants (packed arithmetic, radix set, old screen).
O=gpurun_out
mkdir -p $O
P=ska-sdp-func_b200
( time timeout 900 python -m pytest tests -m gpu -q --durations=8 --deselect tests/test_gpu_timed_shapes.py::test_c4_grid_size_24640_single_line_mode ) > $O/pytest_gpu2.log 2>&1
tail -15 $O/pytest_gpu2.log
timeout 900 python profiles/debug/c4_parity.py > $O/c4_parity.txt 2>&1; tail -5 $O/c4_parity.txt
for v in default nopacked c2set oldnp; do
  for d in grid degrid; do
    if [ $v = default ]; then L=$P/lib; else L=$P/lib_$v; fi
    SKA_SDP_FUNC_LIB_DIR=$L timeout 300 python profiles/run_path.py C2 $d 4 > $O/ab_c2_${d}_$v.txt 2>&1
    echo "C2 $d $v: $(tail -2 $O/ab_c2_${d}_$v.txt | head -1 | cut -c1-330)"
  done
done
for v in default nopacked oldnp; do
  if [ $v = default ]; then L=$P/lib; else L=$P/lib_$v; fi
  SKA_SDP_FUNC_LIB_DIR=$L timeout 400 python profiles/run_path.py C4 grid 2 > $O/ab_c4_grid_$v.txt 2>&1
  echo "C4 grid $v: $(tail -2 $O/ab_c4_grid_$v.txt | head -1 | cut -c1-330)"
  SKA_SDP_FUNC_LIB_DIR=$L SKA_SDP_FUNC_ES_FFT_RADICES=16,14,11,10 SKA_SDP_FUNC_ES_FFT_CFG=1024,1024,0,0 timeout 400 python profiles/run_path.py C4 grid 2 > $O/ab_c4_grid_wide_$v.txt 2>&1
  echo "C4 grid wide $v: $(tail -2 $O/ab_c4_grid_wide_$v.txt | head -1 | cut -c1-330)"
done
