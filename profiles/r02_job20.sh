#!/bin/bash
# GPU-box job 20 (round 2): tiled functions (parity vs oracle and vs the
# reference CUDA build, timings), neighbouring GPU tests, the C3 line.
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_tiled_functions.py -m gpu -q -s > $O/pytest_tiled20.log 2>&1; tail -5 $O/pytest_tiled20.log
for s in 0 1; do
  timeout 600 python profiles/run_tiled.py --single $s > $O/r02_tiled_s$s.json 2> $O/r02_tiled_s$s.err; tail -c 1200 $O/r02_tiled_s$s.json; tail -3 $O/r02_tiled_s$s.err
done
timeout 600 python -m pytest tests/test_weighting.py tests/test_zz_no_alloc.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest_gpu20.log 2>&1; tail -3 $O/pytest_gpu20.log
timeout 900 python bench.py --workload C3 --steps 3 --warmup 3 > $O/r02_bench_C3_n1.json 2> $O/bench_C3.err; tail -c 600 $O/r02_bench_C3_n1.json
