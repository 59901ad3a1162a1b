#!/bin/bash
# GPU-box job 10 (round 2): zero-butterfly skipping in the row kernels; C2 bench
# line with the reference-GPU leg; ncu captures of the C4 kernels.
O=gpurun_out
mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q --durations=5 ) > $O/pytest_gpu10.log 2>&1
tail -7 $O/pytest_gpu10.log
for w in C2 C4 C5; do
  for d in grid degrid; do
    timeout 400 python profiles/run_path.py $w $d 3 > $O/ab10_${w}_${d}.txt 2>&1
    echo "$w $d: $(tail -2 $O/ab10_${w}_${d}.txt | head -1 | grep -o "'ms_total[^}]*'ms_screen': [0-9.]*")"
  done
done
( time timeout 900 python bench.py --workload C2 --steps 5 --warmup 3 ) > $O/r02_b_c2_v2.json 2> $O/r02_b_c2_v2.err; cut -c1-250 $O/r02_b_c2_v2.json; tail -4 $O/r02_b_c2_v2.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:es_fft_grid -s 2 -c 2 -o $O/r02_ncu_c4_grid_c -f python profiles/run_path.py C4 grid 1 > $O/ncu_c4g.log 2>&1; tail -2 $O/ncu_c4g.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:es_fft_degrid -s 2 -c 2 -o $O/r02_ncu_c4_degrid_c -f python profiles/run_path.py C4 degrid 1 > $O/ncu_c4d.log 2>&1; tail -2 $O/ncu_c4d.log
