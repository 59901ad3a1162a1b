#!/bin/bash
# GPU-box job 12 (round 2): evidence for profiles/ -- bench lines of every
# configuration (N = 1), launch lists, ncu --set full captures.
O=gpurun_out
mkdir -p $O
( time timeout 1500 python bench.py ) > $O/r02_bench_C4_n1.json 2> $O/r02_bench_C4_n1.err; cut -c1-200 $O/r02_bench_C4_n1.json; tail -4 $O/r02_bench_C4_n1.err
for w in C2 C3 C5 C1; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-ref-gpu > $O/r02_bench_${w}_n1.json 2> $O/r02_bench_${w}_n1.err; cut -c1-160 $O/r02_bench_${w}_n1.json
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r02_bench_C4_reference.json 2> $O/r02_bench_C4_reference.err; cut -c1-200 $O/r02_bench_C4_reference.json
timeout 600 python bench.py --impl reference-gpu --workload C5 --steps 1 --warmup 1 > $O/r02_bench_C5_reference_gpu.json 2> $O/r02_bench_C5_reference_gpu.err; cut -c1-200 $O/r02_bench_C5_reference_gpu.json
# launch lists (gpu__time_duration only): one C4 gridding call, C2 grid + degrid
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02_launches_c4_grid.csv python profiles/run_path.py C4 grid 1 > $O/ncu_l_c4.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02_launches_c2_grid.csv python profiles/run_path.py C2 grid 2 > $O/ncu_l_c2g.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02_launches_c2_degrid.csv python profiles/run_path.py C2 degrid 2 > $O/ncu_l_c2d.log 2>&1
# full captures: C2 (all hot kernels of one batch), C4 / C5 column + row kernels
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"es_fft_grid|es_scatter2" -s 6 -c 3 -o $O/r02_ncu_c2_grid_final -f python profiles/run_path.py C2 grid 3 > $O/ncu_f_c2g.log 2>&1; tail -1 $O/ncu_f_c2g.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"es_fft_degrid|es_gather2" -s 6 -c 3 -o $O/r02_ncu_c2_degrid_final -f python profiles/run_path.py C2 degrid 3 > $O/ncu_f_c2d.log 2>&1; tail -1 $O/ncu_f_c2d.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"es_fft_grid|es_scatter2" -s 3 -c 3 -o $O/r02_ncu_c4_grid_final -f python profiles/run_path.py C4 grid 1 > $O/ncu_f_c4g.log 2>&1; tail -1 $O/ncu_f_c4g.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"es_fft_grid" -s 4 -c 2 -o $O/r02_ncu_c5_grid_final -f python profiles/run_path.py C5 grid 1 > $O/ncu_f_c5g.log 2>&1; tail -1 $O/ncu_f_c5g.log
